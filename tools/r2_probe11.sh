#!/bin/bash
# r2 probe 11 (2 GPUs): number of z chunks per step in the pipelined scheme now that the x faces are fused
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
O=gpurun_out/p11
port=29520
for nz in 0 0.1; do
  for c in 1 2 3 4 6; do
    port=$((port+1))
    GSB_HALO_CHUNKS=$c python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 2 --steps 20 --warmup 5 --windows 2 --no-e2e --no-parity --no-strong --noise $nz > ${O}_c${c}_nz${nz}.json 2> ${O}_c${c}_nz${nz}.err
    sleep 2
  done
done
echo done
