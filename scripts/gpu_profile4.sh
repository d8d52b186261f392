#!/bin/bash
# ncu --set full of the stand-alone dgrad and wgrad GEMMs at the five conv layer shapes (one launch each)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
GEMM_ONCE=1 timeout 300 python scripts/bench_gemm.py dgrad wgrad > gpurun_out/plain_gemm.log 2>&1 &&
GEMM_ONCE=1 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"gemm" -c 10 -o gpurun_out/prof_gemms python scripts/bench_gemm.py dgrad wgrad > gpurun_out/ncu_gemm.log 2>&1
echo "ncu exit=$?"; tail -3 gpurun_out/ncu_gemm.log
