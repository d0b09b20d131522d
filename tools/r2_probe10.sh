#!/bin/bash
# r2 probe 10: Float64 noise with the Philox block shared by lane pairs across two planes
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p10
timeout 2400 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity2.py -x -q -m gpu -k "Float64 or float64 or f64 or noise or halo or x_faces or restart" > ${O}_pytest.log 2>&1; echo "pytest rc=$?" >> ${O}_pytest.log
T=grayscott.jl_b200/gs_tune
$T 768 f64 ref 1 10 tmalo:4:4:32 tmalo:8:4:64 tmalo:8:6:32 tmalo:4:8:32 tmahi:8:4:64 tmalo:4:4:32 > ${O}_f64_1.log 2>&1
$T 768 f64 ref 0 10 tmalo:8:6:32 > ${O}_f64_0.log 2>&1
sleep 3
$T 768 f64 ref 1 200 tmalo:4:4:32 > ${O}_f64_1_sustained.log 2>&1
echo done
