#!/bin/bash
# r2 probe 7 (2 GPUs): bench.py at N = 2 over NCCL with the x faces fused (default) and strided, the NCCL
# tests, strong scaling of config 4
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p7
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 "${@:2}"; }
run 29511 --steps 20 --warmup 5 > ${O}_bench_n2.json 2> ${O}_bench_n2.err; echo "rc=$?" >> ${O}_bench_n2.err
GSB_FUSE_X=0 run 29512 --steps 20 --warmup 5 --no-e2e --no-parity --no-strong > ${O}_bench_n2_fuse0.json 2> ${O}_bench_n2_fuse0.err
run 29513 --steps 20 --warmup 5 --no-e2e --no-parity --no-strong --noise 0 > ${O}_bench_n2_noise0.json 2> ${O}_bench_n2_noise0.err
run 29514 --steps 30 --warmup 5 --scaling strong --global-L 768 --precision Float64 --no-e2e > ${O}_strong768_n2.json 2> ${O}_strong768_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus 2 --impl reference --steps 5 --warmup 1 > ${O}_bench_n2_ref.json 2> ${O}_bench_n2_ref.err
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -m gpu > ${O}_pytest_multi.log 2>&1; echo "pytest rc=$?" >> ${O}_pytest_multi.log
echo done
