#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q 2>&1 | tail -2
echo "=== default"; timeout 300 python scripts/bench_gemm.py fwd8 dgrad wgrad
echo "=== swap tiles 192"; ZKB_SWAP_BOXES=3 timeout 300 python scripts/bench_gemm.py fwd8 dgrad | head -2
echo "=== again default"; timeout 300 python scripts/bench_gemm.py fwd8 dgrad wgrad
} 2>&1 | tee gpurun_out/probe_order.log
