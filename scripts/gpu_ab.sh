#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
echo "== gemm parity tests"; timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_train_step.py -m gpu -x -q 2>&1 | tail -6
echo "== default (swap for N=128)"; timeout 300 python scripts/bench_gemm.py dgrad
echo "== swap off"; ZKB_KGEMM_SWAP=0 timeout 300 python scripts/bench_gemm.py dgrad
echo "== swap on everywhere"; ZKB_KGEMM_SWAP=1 timeout 300 python scripts/bench_gemm.py dgrad
timeout 900 python bench.py --steps 10 --warmup 3 --breakdown --no-cpu-baseline > gpurun_out/bench1.json 2> gpurun_out/bench_breakdown.txt
cut -c1-200 gpurun_out/bench1.json; grep TOTAL gpurun_out/bench_breakdown.txt | head -12
} 2>&1 | tee gpurun_out/ab.log
