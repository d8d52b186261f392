#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
echo "== pairs: parity tests (kernels)"; ZKB_KGEMM_PAIRS=1 timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q 2>&1 | tail -15
echo "== pairs off"; timeout 300 python scripts/bench_gemm.py fwd8 dgrad
echo "== pairs on"; ZKB_KGEMM_PAIRS=1 timeout 300 python scripts/bench_gemm.py fwd8 dgrad
} 2>&1 | tee gpurun_out/ab.log
