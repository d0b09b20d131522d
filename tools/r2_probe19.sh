#!/bin/bash
# r2 probe 19 (8 GPUs): the scaling runs as the driver launches them (N = 8, 4, 2, 1 on one box), final kernels
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p19
run() { n=$1; port=$2; shift 2; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n "$@"; }
run 8 29551 --steps 20 --warmup 5 > ${O}_n8.json 2> ${O}_n8.err; echo "rc=$?" >> ${O}_n8.err
run 4 29552 --steps 20 --warmup 5 --no-e2e > ${O}_n4.json 2> ${O}_n4.err; echo "rc=$?" >> ${O}_n4.err
run 2 29553 --steps 20 --warmup 5 --no-e2e > ${O}_n2.json 2> ${O}_n2.err
python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu-baseline > ${O}_n1.json 2> ${O}_n1.err
echo done
