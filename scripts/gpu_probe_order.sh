#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q 2>&1 | tail -2
echo "=== all GEMMs"; timeout 300 python scripts/bench_gemm.py fwd8 dgrad wgrad
echo "=== again"; timeout 300 python scripts/bench_gemm.py fwd8 dgrad wgrad
} 2>&1 | tee gpurun_out/probe_order.log
