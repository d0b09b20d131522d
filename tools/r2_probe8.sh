#!/bin/bash
# r2 probe 8: full GPU suite after the staging / default-shape / native-arithmetic changes; burst and
# sustained timings of the headline kernel; ncu captures for profiles/
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p8
timeout 2400 python -m pytest tests -x -q -m gpu > ${O}_pytest.log 2>&1; echo "pytest rc=$?" >> ${O}_pytest.log
T=grayscott.jl_b200/gs_tune
$T 1024 f32 native 1 10 tmahi:8:6:32 tmahi:8:6:64 tmahi:8:4:32 tmahi:8:6:32 > ${O}_nat1.log 2>&1
$T 1024 f32 native 0 10 tmahi:8:4:16 tmahi:8:4:32 > ${O}_nat0.log 2>&1
sleep 3
$T 1024 f32 native 1 300 tmahi:8:6:32 > ${O}_sustained_nat1.log 2>&1
sleep 3
GSB_TMA_L2PROMO=64 $T 1024 f32 native 1 300 tmahi:8:6:32 > ${O}_sustained_nat1_promo64.log 2>&1
sleep 3
python bench.py --steps 20 --warmup 5 > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?" >> ${O}_bench.err
ncu --set full --import-source on --clock-control none -k regex:stencil --launch-skip 3 -c 1 -f -o ${O}_native_noise_L1024 $T 1024 f32 native 1 2 tmahi:8:6:32 > /dev/null 2>&1
echo done
