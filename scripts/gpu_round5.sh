#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 --breakdown --no-cpu-baseline > gpurun_out/bench1.json 2> gpurun_out/bench_breakdown.txt
echo "bench exit=$?"; cut -c1-200 gpurun_out/bench1.json; grep TOTAL gpurun_out/bench_breakdown.txt | head -14
NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench2.json 2> gpurun_out/bench2.err
echo "bench2 exit=$?"; cat gpurun_out/bench2.json | cut -c1-400; tail -5 gpurun_out/bench2.err
NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --sync-bn > gpurun_out/bench2_syncbn.json 2> gpurun_out/bench2s.err
echo "bench2 syncbn exit=$?"; cat gpurun_out/bench2_syncbn.json | cut -c1-300; tail -3 gpurun_out/bench2s.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 scripts/dp_parity.py 2>&1 | tail -8
