#!/bin/bash
# r2 probe 1: DRAM traffic vs z-chunk length for the headline kernel
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
T=grayscott.jl_b200/gs_tune
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/p1_gpu.txt
$T 1024 f32 native 1 10 tmahi:8:6:16 tmahi:8:6:32 tmahi:8:6:64 tmahi:8:6:128 tmahi:8:4:32 tmahi:8:4:64 tmahi:8:6:32 > gpurun_out/p1_plain_noise.log 2>&1
$T 1024 f32 native 0 10 tmahi:8:4:16 tmahi:8:4:32 tmahi:8:4:64 tmahi:8:4:128 tmahi:8:6:32 tmahi:8:4:16 > gpurun_out/p1_plain_nonoise.log 2>&1
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,lts__t_sectors_srcunit_tex_op_read.sum
for v in tmahi:8:6:16 tmahi:8:6:32 tmahi:8:6:64 tmahi:8:6:128; do
  ncu --metrics $M --clock-control none -k regex:stencil --launch-skip 3 -c 1 --csv --log-file gpurun_out/p1_ncu_noise_${v//:/_}.csv $T 1024 f32 native 1 2 $v > /dev/null 2>&1
done
for p in 0 64; do
  GSB_TMA_L2PROMO=$p ncu --metrics $M --clock-control none -k regex:stencil --launch-skip 3 -c 1 --csv --log-file gpurun_out/p1_ncu_noise_promo${p}.csv $T 1024 f32 native 1 2 tmahi:8:6:32 > /dev/null 2>&1
  GSB_TMA_L2PROMO=$p $T 1024 f32 native 1 10 tmahi:8:6:32 tmahi:8:6:64 > gpurun_out/p1_plain_promo${p}.log 2>&1
done
echo done
