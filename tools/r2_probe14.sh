#!/bin/bash
# r2 probe 14: Float64 divide-by-6 with the rare case behind a call; full suite; L = 2048 on one GPU
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p14
T=grayscott.jl_b200/gs_tune
$T 768 f64 ref 1 10 tmalo:4:4:32 tmalo:8:4:64 tmalo:8:6:32 tmalo:4:4:32 > ${O}_f64_1.log 2>&1
$T 768 f64 ref 0 10 tmalo:8:6:32 tmalo:4:8:32 > ${O}_f64_0.log 2>&1
$T 256 f64 ref 1 50 tmalo:4:4:16 > ${O}_f64_L256.log 2>&1
timeout 2400 python -m pytest tests -x -q -m gpu > ${O}_pytest.log 2>&1; echo "pytest rc=$?" >> ${O}_pytest.log
python bench.py --scaling strong --global-L 2048 --steps 5 --warmup 3 --windows 1 --no-e2e --no-cpu-baseline --no-ref-arith > ${O}_L2048_one_gpu.json 2> ${O}_L2048_one_gpu.err
echo done
