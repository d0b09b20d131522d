#!/bin/bash
# r2 probe 2: instruction-count reductions (mbarrier suspend hint, Philox round keys from the host,
# sign-transfer in the exact divide-by-6) against the round-1 kernels (build/exp/base)
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p2
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "kat or bit_identical or one_step or variants or noise or config2" > ${O}_pytest.log 2>&1
echo "pytest rc=$?" >> ${O}_pytest.log
for T in build/exp/base/gs_tune grayscott.jl_b200/gs_tune; do
  tag=$(basename $(dirname $T))
  $T 1024 f32 native 1 10 tmahi:8:6:32 tmahi:8:6:64 tmahi:8:4:32 tmahi:8:6:32 > ${O}_${tag}_nat1.log 2>&1
  $T 1024 f32 native 0 10 tmahi:8:4:16 tmahi:8:4:32 tmahi:8:6:32 tmahi:8:4:16 > ${O}_${tag}_nat0.log 2>&1
  $T 1024 f32 ref 0 10 tmahi:8:4:64 tmahi:8:6:32 tmalo:4:8:64 tmahi:8:4:64 > ${O}_${tag}_ref0.log 2>&1
  $T 1024 f32 ref 1 10 tmahi:8:4:64 tmalo:4:8:64 tmalo:4:4:32 tmahi:8:6:32 > ${O}_${tag}_ref1.log 2>&1
  $T 768 f64 ref 0 10 tmalo:8:6:32 tmahi:8:6:64 tmalo:4:8:32 > ${O}_${tag}_f64_0.log 2>&1
  $T 768 f64 ref 1 10 tmalo:8:4:64 tmahi:8:4:64 tmalo:4:4:32 tmalo:8:6:32 > ${O}_${tag}_f64_1.log 2>&1
done
echo done
