#!/bin/bash
# r2 probe 9: full GPU suite; sustained timing of the headline kernel (elected-lane slot release vs the
# round-1 form); bench at N = 1 (driver's flags)
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p9
timeout 2400 python -m pytest tests -x -q -m gpu > ${O}_pytest.log 2>&1; echo "pytest rc=$?" >> ${O}_pytest.log
for d in grayscott.jl_b200 build/exp/noelect; do
  tag=$(basename $d)
  sleep 3
  $d/gs_tune 1024 f32 native 1 300 tmahi:8:6:32 > ${O}_${tag}_sustained_nat1.log 2>&1
  sleep 3
  $d/gs_tune 1024 f32 native 1 10 tmahi:8:6:32 tmahi:8:6:32 > ${O}_${tag}_short_nat1.log 2>&1
done
sleep 3
python bench.py --steps 20 --warmup 5 > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?" >> ${O}_bench.err
echo done
