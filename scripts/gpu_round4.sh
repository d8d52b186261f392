#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 --breakdown > gpurun_out/bench.json 2> gpurun_out/bench_breakdown.txt
echo "bench exit=$?"; cat gpurun_out/bench.json | cut -c1-300; head -30 gpurun_out/bench_breakdown.txt; grep TOTAL gpurun_out/bench_breakdown.txt | head -12
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 94 -c 2 -o gpurun_out/prof_r1_bwd1 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu exit=$?"
