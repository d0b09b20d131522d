#!/bin/bash
# r2 probe 6: edge-tile-only x-face launches; reference-arithmetic variants chosen by a uniform branch;
# sustained (300-step) against short (10-step) timings, with clocks
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p6
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity2.py -x -q -m gpu -k "x_faces or bit_identical or one_step or config2 or kat or restart or noise_bit" > ${O}_pytest.log 2>&1; echo "pytest rc=$?" >> ${O}_pytest.log
T=grayscott.jl_b200/gs_tune
for fx in 2 1 0; do
  GSB_FUSE_X=$fx GS_TUNE_RANKS=2 timeout 300 $T 1024 f32 native 1 10 tmahi:8:6:32 > ${O}_r2_fuse${fx}_nat1.log 2>&1
  GSB_FUSE_X=$fx GS_TUNE_RANKS=2 timeout 300 $T 1024 f32 ref 1 10 tmahi:8:4:64 > ${O}_r2_fuse${fx}_ref1.log 2>&1
  GSB_FUSE_X=$fx GS_TUNE_RANKS=2 timeout 300 $T 768 f64 ref 1 10 tmalo:8:4:64 > ${O}_r2_fuse${fx}_f64.log 2>&1
  GSB_FUSE_X=$fx GS_TUNE_RANKS=8 timeout 300 $T 512 f32 native 1 10 tmahi:8:6:32 > ${O}_r8_fuse${fx}_nat1.log 2>&1
done
$T 1024 f32 ref 0 10 tmahi:8:4:64 tmalo:4:8:64 tmahi:8:6:32 > ${O}_ref0.log 2>&1
$T 1024 f32 ref 1 10 tmahi:8:4:64 tmalo:4:8:64 tmalo:8:4:64 tmahi:8:6:32 > ${O}_ref1.log 2>&1
$T 768 f64 ref 0 10 tmalo:8:6:32 tmalo:4:8:32 > ${O}_f64_0.log 2>&1
$T 768 f64 ref 1 10 tmalo:8:4:64 tmalo:4:4:32 tmalo:8:6:32 tmahi:8:4:64 > ${O}_f64_1.log 2>&1
# sustained: clocks sampled while 300 steps run
(nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_throttle_reasons.sw_power_cap --format=csv,noheader -lms 100 > ${O}_smi.log 2>&1 &) 
$T 1024 f32 native 1 300 tmahi:8:6:32 > ${O}_sustained_nat1.log 2>&1
sleep 2
$T 1024 f32 native 0 300 tmahi:8:4:16 > ${O}_sustained_nat0.log 2>&1
sleep 2
$T 1024 f32 native 1 10 tmahi:8:6:32 > ${O}_short_nat1.log 2>&1
pkill -x nvidia-smi
echo done
