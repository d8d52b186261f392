#!/bin/bash
# configs[0] shape (batch 128) and batch 1024 on one GPU: per-kernel launches vs CUDA-graph replay
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
pr() { python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['config'].get('launch'))"; }
{
for b in 128 1024; do
echo -n "batch $b per-kernel: "; timeout 600 python bench.py --steps 50 --warmup 10 --batch $b --no-cpu-baseline 2>/dev/null | pr
echo -n "batch $b graph:      "; timeout 600 python bench.py --steps 50 --warmup 10 --batch $b --no-cpu-baseline --graph 2>/dev/null | pr
done
} | tee gpurun_out/small_batch.log
