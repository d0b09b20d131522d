#!/bin/bash
# r2 probe 20: what the driver runs at round end, with the final code: GPU test suite, smoke, bench (own arm and
# reference arm), plus the ncu evidence of the final kernels (compact text: the reports are ~30 MB each)
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p20
T=grayscott.jl_b200/gs_tune
timeout 2400 python -m pytest tests -x -q -m gpu > ${O}_pytest.log 2>&1; echo "pytest rc=$?" >> ${O}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > ${O}_smoke.log 2>&1; echo "smoke rc=$?" >> ${O}_smoke.log
sleep 5
python bench.py --steps 20 --warmup 5 > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?" >> ${O}_bench.err
python bench.py --impl reference --steps 20 --warmup 5 > ${O}_bench_ref.json 2> ${O}_bench_ref.err
B="python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-ref-arith --windows 1"
$B > ${O}_plain_bench.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 5 -c 20 --csv --log-file ${O}_launches_bench.csv $B > ${O}_ncu_bench.log 2>&1
GS_TUNE_RANKS=2 $T 512 f32 native 1 4 tmahi:8:6:32 > ${O}_plain_two.log 2>&1 && GS_TUNE_RANKS=2 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 24 -c 32 --csv --log-file ${O}_launches_two_subdomains.csv $T 512 f32 native 1 4 tmahi:8:6:32 > ${O}_ncu_two.log 2>&1
cap() { name=$1; iters=$2; shift 2
  ncu --set full --import-source on --clock-control none -k regex:stencil -s 3 -c 1 -f -o ${O}_$name $T "$@" > /dev/null 2>&1
  ncu -i ${O}_$name.ncu-rep --page raw --csv > ${O}_$name.raw.csv 2>/dev/null
  ncu -i ${O}_$name.ncu-rep --page source --csv --print-source sass 2>/dev/null | python tools/ncu_dyn_mix.py $iters > ${O}_$name.mix.txt 2>&1
  rm -f ${O}_$name.ncu-rep
}
cap native_noise_L1024 $((1024*1024*1024/128)) 1024 f32 native 1 2 tmahi:8:6:32
cap ref_noise_L1024 $((1024*1024*1024/128)) 1024 f32 ref 1 2 tmahi:8:4:64
cap f64_noise_L768 $((768*768*768/64)) 768 f64 ref 1 2 tmalo:4:4:32
cap native_nonoise_L1024 $((1024*1024*1024/128)) 1024 f32 native 0 2 tmahi:8:4:16
# ring depth of the headline kernel: 20 steps (what bench.py times) and 300 steps
for v in tmahi:8:6:32 tmahi:8:4:32; do sleep 4; $T 1024 f32 native 1 20 $v >> ${O}_burst20.log 2>&1; done
for v in tmahi:8:6:32 tmahi:8:4:32; do sleep 4; $T 1024 f32 native 1 300 $v >> ${O}_sustained300.log 2>&1; done
echo done
