#!/bin/bash
# ncu --set full of the forward elementwise kernels (first conv + BatchNorm + sign; max-pool; BatchNorm apply)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"conv_first_bn_sign|maxpool2_i16" -s 8 -c 2 -o gpurun_out/prof_elem python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit=$?"
