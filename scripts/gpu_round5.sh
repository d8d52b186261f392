#!/bin/bash
# 2-GPU visit: DP-vs-single-GPU parity, N=2 bench (per-replica BN and SyncBN), all-reduce bandwidth
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -15
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench2.json 2> gpurun_out/bench2.err
echo "bench2 exit=$?"; cat gpurun_out/bench2.json | cut -c1-200
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --sync-bn > gpurun_out/bench2_syncbn.json 2> gpurun_out/bench2s.err
echo "bench2 syncbn exit=$?"; cat gpurun_out/bench2_syncbn.json | cut -c1-200
