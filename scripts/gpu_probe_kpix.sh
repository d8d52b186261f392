#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
for k in 64 48 32; do
echo "=== wgrad kpix=$k"; ZKB_WGRAD_KPIX=$k timeout 300 python scripts/bench_gemm.py wgrad
done
echo "=== wgrad kpix=48 waves=1"; ZKB_WGRAD_KPIX=48 ZKB_WGRAD_WAVES=1 timeout 300 python scripts/bench_gemm.py wgrad
echo "=== wgrad kpix=64 waves=1"; ZKB_WGRAD_WAVES=1 timeout 300 python scripts/bench_gemm.py wgrad
echo "=== wgrad kpix=64 tpi=1"; ZKB_WGRAD_TPI=1 timeout 300 python scripts/bench_gemm.py wgrad
} 2>&1 | tee gpurun_out/probe_kpix.log
