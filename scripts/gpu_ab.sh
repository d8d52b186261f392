#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
for w in 2 3 4 8; do
echo "== reduce waves=$w"; ZKB_REDUCE_WAVES=$w timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --breakdown 2> gpurun_out/bd_$w.txt | cut -c1-150
grep -E "TOTAL zkb_(bn_bwd_reduce|bn_stats|maxpool)" gpurun_out/bd_$w.txt
done
} 2>&1 | tee gpurun_out/ab.log
