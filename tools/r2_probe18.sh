#!/bin/bash
# r2 probe 18: partial-vector stores behind a call (no second pointer copy in the hot loop); x faces fused in every
# CTA against edge tiles only, now that the XF instantiation no longer spills
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p18
T=grayscott.jl_b200/gs_tune
timeout 2400 python -m pytest tests -x -q -m gpu > ${O}_pytest.log 2>&1; echo "pytest rc=$?" >> ${O}_pytest.log
$T 1024 f32 native 1 10 tmahi:8:6:32 tmahi:8:4:32 tmahi:8:6:32 > ${O}_nat1.log 2>&1
$T 1024 f32 native 0 10 tmahi:8:4:16 > ${O}_nat0.log 2>&1
$T 1024 f32 ref 1 10 tmahi:8:4:64 > ${O}_ref1.log 2>&1
$T 1024 f32 ref 0 10 tmahi:8:4:64 > ${O}_ref0.log 2>&1
$T 768 f64 ref 1 10 tmalo:4:4:32 > ${O}_f64_1.log 2>&1
$T 768 f64 ref 0 10 tmalo:8:6:32 > ${O}_f64_0.log 2>&1
sleep 3
$T 1024 f32 native 1 300 tmahi:8:6:32 > ${O}_sustained_nat1.log 2>&1
sleep 3
for fx in 2 1 0; do
  GSB_FUSE_X=$fx GS_TUNE_RANKS=2 timeout 300 $T 1024 f32 native 1 10 tmahi:8:6:32 > ${O}_r2_fuse${fx}_nat1.log 2>&1
  GSB_FUSE_X=$fx GS_TUNE_RANKS=2 timeout 300 $T 768 f64 ref 1 10 tmalo:4:4:32 > ${O}_r2_fuse${fx}_f64.log 2>&1
done
echo done
