#!/bin/bash
# r2 probe 23: the GPU suite and smoke on the final tree
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/p23_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/p23_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/p23_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/p23_smoke.log
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/p23_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/p23_noalloc_alone.log
echo done
