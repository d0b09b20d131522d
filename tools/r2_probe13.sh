#!/bin/bash
# r2 probe 13: evidence for profiles/ -- launch list of the bench command, launch list of a two-subdomain
# pipelined step (x faces inside the update kernel), ncu --set full of the three slowest-to-roofline variants,
# TMA L2 promotion 64 B against 128 B
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p13
T=grayscott.jl_b200/gs_tune
B="python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-ref-arith --windows 1"
$B > ${O}_plain_bench.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 5 -c 20 --csv --log-file ${O}_launches_bench.csv $B > ${O}_ncu_bench.log 2>&1
GS_TUNE_RANKS=2 $T 512 f32 native 1 4 tmahi:8:6:32 > ${O}_plain_two.log 2>&1 && GS_TUNE_RANKS=2 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 40 -c 60 --csv --log-file ${O}_launches_two_subdomains.csv $T 512 f32 native 1 4 tmahi:8:6:32 > ${O}_ncu_two.log 2>&1
# the reports are ~30 MB each and gpurun_out/ is capped at 64 MiB: keep the raw page, the per-opcode dynamic mix
# and the stall samples as text, drop the report
cap() { name=$1; iters=$2; shift 2
  ncu --set full --import-source on --clock-control none -k regex:stencil -s 3 -c 1 -f -o ${O}_$name $T "$@" > /dev/null 2>&1
  ncu -i ${O}_$name.ncu-rep --page raw --csv > ${O}_$name.raw.csv 2>/dev/null
  ncu -i ${O}_$name.ncu-rep --page source --csv --print-source sass 2>/dev/null | python tools/ncu_dyn_mix.py $iters > ${O}_$name.mix.txt 2>&1
  rm -f ${O}_$name.ncu-rep
}
cap native_noise_L1024 $((1024*1024*1024/128)) 1024 f32 native 1 2 tmahi:8:6:32
cap ref_noise_L1024 $((1024*1024*1024/128)) 1024 f32 ref 1 2 tmalo:8:4:64
cap f64_noise_L768 $((768*768*768/64)) 768 f64 ref 1 2 tmalo:4:4:32
cap native_nonoise_L1024 $((1024*1024*1024/128)) 1024 f32 native 0 2 tmahi:8:4:16
for p in 128 64; do
  GSB_TMA_L2PROMO=$p $T 1024 f32 native 1 10 tmahi:8:6:32 tmahi:8:6:32 > ${O}_promo${p}_nat1.log 2>&1
  GSB_TMA_L2PROMO=$p $T 1024 f32 native 0 10 tmahi:8:4:16 tmahi:8:4:16 > ${O}_promo${p}_nat0.log 2>&1
  GSB_TMA_L2PROMO=$p $T 768 f64 ref 0 10 tmalo:8:6:32 tmalo:8:6:32 > ${O}_promo${p}_f64_0.log 2>&1
  GSB_TMA_L2PROMO=$p $T 768 f64 ref 1 10 tmalo:4:4:32 > ${O}_promo${p}_f64_1.log 2>&1
  GSB_TMA_L2PROMO=$p $T 1024 f32 ref 0 10 tmalo:4:8:64 > ${O}_promo${p}_ref0.log 2>&1
  GSB_TMA_L2PROMO=$p $T 256 f32 native 1 50 tmahi:4:4:16 > ${O}_promo${p}_L256.log 2>&1
done
echo done
