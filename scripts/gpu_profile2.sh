#!/bin/bash
# one full ncu capture of the two dominant GEMMs of a step: layer 1's dgrad with the fused first-layer
# backward (kgemm) and layer 1's wgrad (wgemm) = launches 7 and 8 of the 9 matching launches per step
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"kgemm_first_bf16|wgemm_bf16" -s 34 -c 2 -o gpurun_out/prof_bwd1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit=$?"
