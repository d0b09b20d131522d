#!/bin/bash
# r2 probe 22: the headline default with 4 ring slots: bench line, launch list, ncu capture, the tests that name the shape
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p22
T=grayscott.jl_b200/gs_tune
timeout 1200 python -m pytest tests/test_gpu_parity2.py -x -q -m gpu -k "headline or native or full_size" > ${O}_pytest.log 2>&1; echo "pytest rc=$?" >> ${O}_pytest.log
sleep 5
python bench.py --steps 20 --warmup 5 > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?" >> ${O}_bench.err
B="python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-ref-arith --windows 1"
$B > ${O}_plain_bench.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 5 -c 20 --csv --log-file ${O}_launches_bench.csv $B > ${O}_ncu_bench.log 2>&1
cap() { name=$1; iters=$2; shift 2
  ncu --set full --import-source on --clock-control none -k regex:stencil -s 3 -c 1 -f -o ${O}_$name $T "$@" > /dev/null 2>&1
  ncu -i ${O}_$name.ncu-rep --page raw --csv > ${O}_$name.raw.csv 2>/dev/null
  ncu -i ${O}_$name.ncu-rep --page source --csv --print-source sass 2>/dev/null | python tools/ncu_dyn_mix.py $iters > ${O}_$name.mix.txt 2>&1
  rm -f ${O}_$name.ncu-rep
}
cap native_noise_L1024 $((1024*1024*1024/128)) 1024 f32 native 1 2 tmahi:8:4:32
echo done
