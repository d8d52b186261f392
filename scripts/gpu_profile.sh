#!/bin/bash
# launch list (shares) + one full ncu capture of the backward GEMMs of layer 1
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 273 -c 100 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
echo "launch list exit=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 94 -c 2 -o gpurun_out/prof_bwd1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit=$?"
