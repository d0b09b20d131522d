#!/bin/bash
# r2 probe 3: pipe throughputs; Float64 kernels after restoring the exact divide-by-6 slow-path test
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p3
build/tools/ubench_pipes > ${O}_pipes.log 2>&1
T=grayscott.jl_b200/gs_tune
$T 768 f64 ref 0 10 tmalo:8:6:32 tmahi:8:6:64 tmalo:4:8:32 > ${O}_f64_0.log 2>&1
$T 768 f64 ref 1 10 tmalo:8:4:64 tmahi:8:4:64 tmalo:4:4:32 tmalo:8:6:32 > ${O}_f64_1.log 2>&1
$T 1024 f32 ref 1 10 tmahi:8:4:64 tmalo:4:8:64 tmalo:8:4:64 tmalo:8:6:32 tmahi:4:4:32 > ${O}_ref1.log 2>&1
echo done
