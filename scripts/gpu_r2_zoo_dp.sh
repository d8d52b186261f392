#!/bin/bash
# r2 zoo models, data-parallel (N GPUs = $1, default 2): both bench lines, weak scaling
cd "$(dirname "$0")/.."
N=${1:-2}
mkdir -p gpurun_out
for m in alexnet bireal18; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
  scripts/bench_zoo.py --model $m --steps 10 --warmup 3 > gpurun_out/zoo_${m}_n$N.json 2> gpurun_out/zoo_${m}_n$N.err
echo "$m N=$N exit=$?"; cat gpurun_out/zoo_${m}_n$N.json; tail -3 gpurun_out/zoo_${m}_n$N.err
done
