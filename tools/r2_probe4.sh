#!/bin/bash
# r2 probe 4: whole GPU test suite, smoke, bench at N = 1, pipe throughputs, ncu source capture of the
# Float32 reference-arithmetic kernel with noise
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p4
python -m pytest tests -x -q -m gpu > ${O}_pytest.log 2>&1; echo "pytest rc=$?" >> ${O}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > ${O}_smoke.log 2>&1; echo "smoke rc=$?" >> ${O}_smoke.log
python bench.py > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?" >> ${O}_bench.err
python bench.py --impl reference > ${O}_bench_ref.json 2> ${O}_bench_ref.err
build/tools/ubench_pipes > ${O}_pipes.log 2>&1
T=grayscott.jl_b200/gs_tune
ncu --set full --import-source on --clock-control none -k regex:stencil --launch-skip 3 -c 1 -f -o ${O}_ref_noise_L1024 $T 1024 f32 ref 1 2 tmahi:8:4:64 > /dev/null 2>&1
ncu --set full --import-source on --clock-control none -k regex:stencil --launch-skip 3 -c 1 -f -o ${O}_f64_noise_L768 $T 768 f64 ref 1 2 tmalo:8:4:64 > /dev/null 2>&1
echo done
