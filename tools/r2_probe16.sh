#!/bin/bash
# r2 probe 16: one running store pointer; dynamic instruction mix of the headline kernel after the loop changes
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p16
T=grayscott.jl_b200/gs_tune
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity2.py -x -q -m gpu -k "one_step or variants or kat or noise_bit or x_faces or full_size or tiny" > ${O}_pytest.log 2>&1; echo "pytest rc=$?" >> ${O}_pytest.log
$T 1024 f32 native 1 10 tmahi:8:6:32 tmahi:8:4:32 tmahi:8:6:32 > ${O}_nat1.log 2>&1
$T 1024 f32 native 0 10 tmahi:8:4:16 tmahi:8:4:32 > ${O}_nat0.log 2>&1
$T 1024 f32 ref 1 10 tmahi:8:4:64 > ${O}_ref1.log 2>&1
$T 768 f64 ref 1 10 tmalo:4:4:32 > ${O}_f64_1.log 2>&1
sleep 3
$T 1024 f32 native 1 300 tmahi:8:6:32 > ${O}_sustained_nat1.log 2>&1
sleep 3
$T 1024 f32 native 1 300 tmahi:8:4:32 > ${O}_sustained_nat1_s4.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:stencil -s 3 -c 1 -f -o ${O}_nn $T 1024 f32 native 1 2 tmahi:8:6:32 > /dev/null 2>&1
ncu -i ${O}_nn.ncu-rep --page raw --csv > ${O}_native_noise_L1024.raw.csv 2>/dev/null
ncu -i ${O}_nn.ncu-rep --page source --csv --print-source sass 2>/dev/null | python tools/ncu_dyn_mix.py $((1024*1024*1024/128)) > ${O}_native_noise_L1024.mix.txt 2>&1
rm -f ${O}_nn.ncu-rep
echo done
