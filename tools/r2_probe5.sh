#!/bin/bash
# r2 probe 5: x faces inside the update kernel (producer-warp bulk copies) against the strided pack / unpack
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p5
timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "x_faces or halo_exchange_bit_exact or overlap or one_step" > ${O}_pytest.log 2>&1; echo "pytest rc=$?" >> ${O}_pytest.log
T=grayscott.jl_b200/gs_tune
for fx in 1 0; do
  GSB_FUSE_X=$fx GS_TUNE_RANKS=2 timeout 300 $T 1024 f32 native 1 10 tmahi:8:6:32 > ${O}_r2_fuse${fx}_nat1.log 2>&1
  GSB_FUSE_X=$fx GS_TUNE_RANKS=2 timeout 300 $T 1024 f32 native 0 10 tmahi:8:4:16 > ${O}_r2_fuse${fx}_nat0.log 2>&1
  GSB_FUSE_X=$fx GS_TUNE_RANKS=2 timeout 300 $T 768 f64 ref 1 10 tmalo:8:4:64 > ${O}_r2_fuse${fx}_f64.log 2>&1
done
echo done
