#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
echo "=== fwd8/dgrad default (pairs only for two-term)"; timeout 300 python scripts/bench_gemm.py fwd8 dgrad
echo "=== pairs forced on"; ZKB_KGEMM_PAIRS=1 timeout 300 python scripts/bench_gemm.py fwd8 dgrad
echo "=== pairs forced off"; ZKB_KGEMM_PAIRS=0 timeout 300 python scripts/bench_gemm.py fwd8 dgrad
echo "=== fwd8 loads only"; ZKB_GEMM_DBG=1 timeout 300 python scripts/bench_gemm.py fwd8
echo "=== fwd8 MMAs only"; ZKB_GEMM_DBG=2 timeout 300 python scripts/bench_gemm.py fwd8
echo "=== fwd8 loads only, pairs"; ZKB_KGEMM_PAIRS=1 ZKB_GEMM_DBG=1 timeout 300 python scripts/bench_gemm.py fwd8
echo "=== fwd8 MMAs only, pairs"; ZKB_KGEMM_PAIRS=1 ZKB_GEMM_DBG=2 timeout 300 python scripts/bench_gemm.py fwd8
} 2>&1 | tee gpurun_out/probe_pairs.log
