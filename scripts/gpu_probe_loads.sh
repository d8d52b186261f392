#!/bin/bash
# How much of each backward GEMM's time do its TMA loads alone need?  (ZKB_GEMM_DBG=1: no MMA issue)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
echo "=== full"; timeout 300 python scripts/bench_gemm.py dgrad wgrad
echo "=== loads only (no MMA issue)"; ZKB_GEMM_DBG=1 timeout 300 python scripts/bench_gemm.py dgrad wgrad
echo "=== MMAs only (no TMA loads)"; ZKB_GEMM_DBG=2 timeout 300 python scripts/bench_gemm.py dgrad wgrad
} 2>&1 | tee gpurun_out/probe_loads.log
