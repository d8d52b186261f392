#!/bin/bash
# scaling run: N GPUs of one box
N=${1:-8}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench N=$N exit=$?"; cut -c1-220 gpurun_out/bench_n$N.json; tail -3 gpurun_out/bench_n$N.err
