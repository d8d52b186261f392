#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
echo "=== wgrad no merge, full groups first"; ZKB_WGRAD_NO_MERGE=1 timeout 300 python scripts/bench_gemm.py wgrad
echo "=== wgrad interleaved order"; ZKB_WGRAD_ORDER=1 timeout 300 python scripts/bench_gemm.py wgrad
echo "=== wgrad no merge, full groups first, waves 3"; ZKB_WGRAD_WAVES=3 ZKB_WGRAD_NO_MERGE=1 timeout 300 python scripts/bench_gemm.py wgrad
echo "=== wgrad interleaved order waves 3"; ZKB_WGRAD_WAVES=3 ZKB_WGRAD_ORDER=1 timeout 300 python scripts/bench_gemm.py wgrad
echo "=== wgrad interleaved order waves 1"; ZKB_WGRAD_WAVES=1 ZKB_WGRAD_ORDER=1 timeout 300 python scripts/bench_gemm.py wgrad
} 2>&1 | tee gpurun_out/probe_merge.log
