#!/bin/bash
# r2 probe 17: shape sweep after the consumer-loop changes (defaults are re-chosen from this)
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p17
T=grayscott.jl_b200/gs_tune
$T 1024 f32 native 1 10 tmahi:8:6:32 tmahi:8:4:32 tmahi:8:3:32 tmahi:8:4:16 tmahi:8:4:64 tmahi:8:6:16 tmahi:8:6:64 tmahi:4:4:32 tmahi:4:8:32 tmahi:8:4:32 tmahi:8:6:32 > ${O}_nat1.log 2>&1
$T 1024 f32 native 0 10 tmahi:8:4:16 tmahi:8:4:32 tmahi:8:3:32 tmahi:8:6:32 tmahi:4:4:16 tmahi:8:4:16 > ${O}_nat0.log 2>&1
$T 1024 f32 ref 1 10 tmahi:8:4:64 tmahi:8:4:32 tmahi:8:6:32 tmahi:8:3:64 tmalo:8:4:64 tmahi:4:4:32 > ${O}_ref1.log 2>&1
$T 1024 f32 ref 0 10 tmahi:8:4:64 tmahi:8:4:32 tmahi:8:6:32 tmahi:8:4:16 > ${O}_ref0.log 2>&1
$T 768 f64 ref 1 10 tmalo:4:4:32 tmalo:4:4:16 tmalo:4:8:32 tmalo:8:4:32 tmalo:8:6:32 tmahi:4:4:32 tmahi:8:4:32 > ${O}_f64_1.log 2>&1
$T 768 f64 ref 0 10 tmalo:8:6:32 tmalo:8:4:32 tmalo:4:4:32 tmahi:8:4:32 tmahi:8:6:32 > ${O}_f64_0.log 2>&1
sleep 3
for v in tmahi:8:4:32 tmahi:8:6:32 tmahi:8:4:16; do $T 1024 f32 native 1 300 $v >> ${O}_sustained_nat1.log 2>&1; sleep 3; done
echo done
