#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
echo "== gemm parity tests"; timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q 2>&1 | tail -8
echo "== default (tpi=3,kpix=64)"; timeout 300 python scripts/bench_gemm.py
echo "== wgrad tpi=3 kpix=48"; ZKB_WGRAD_KPIX=48 timeout 300 python scripts/bench_gemm.py wgrad
echo "== wgrad tpi=1 kpix=64 (old)"; ZKB_WGRAD_TPI=1 ZKB_WGRAD_KPIX=64 timeout 300 python scripts/bench_gemm.py wgrad
} 2>&1 | tee gpurun_out/ab.log
