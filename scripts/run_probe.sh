#!/bin/bash
# Runs every probe case in its own process with a timeout; summary at the end.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for c in $(python scripts/probe_gemm.py --list); do
  timeout 120 python scripts/probe_gemm.py $c 2>&1 | tail -12
  echo "== $c exit=${PIPESTATUS[0]}"
done 2>&1 | tee gpurun_out/probe_gemm.log
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv | tee -a gpurun_out/probe_gemm.log
