#!/bin/bash
# r2 probe 26: integer -> double of the noise variate by bit pattern + one DADD instead of I2F.F64
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p26
T=grayscott.jl_b200/gs_tune
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity2.py -x -q -m gpu -k "noise or restart or x_faces or halo_exchange_bit" > ${O}_pytest.log 2>&1; echo "pytest rc=$?" >> ${O}_pytest.log
$T 1024 f32 ref 1 10 tmahi:8:4:64 tmahi:8:4:64 > ${O}_ref1.log 2>&1
$T 768 f64 ref 1 10 tmalo:4:4:32 tmalo:4:4:32 > ${O}_f64_1.log 2>&1
echo done
