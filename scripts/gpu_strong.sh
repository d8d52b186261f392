#!/bin/bash
# strong scaling: global batch 4096 split over N GPUs
N=${1:-2}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B=$((4096 / N))
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29545 bench.py --gpus $N --steps 30 --warmup 5 --batch $B --no-cpu-baseline > gpurun_out/bench_strong_n$N.json 2> gpurun_out/bench_strong_n$N.err
echo "strong N=$N exit=$?"; python -c "import json; d=json.load(open('gpurun_out/bench_strong_n$N.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['config'])"
