#!/bin/bash
# r2 widening (1 GPU): 64-channel / 5x5 binary convs, float stems, whole Bi-Real-Net-18 step; then its bench line
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_stem_and_narrow.py tests/test_gpu_bireal.py -m gpu -q -p no:cacheprovider \
  --timeout 240 -o timeout_method=thread -rf 2>&1 | tail -80 | tee gpurun_out/widen_tests.log
timeout 400 python scripts/bench_bireal18.py --breakdown > gpurun_out/bireal18.txt 2> gpurun_out/bireal18.err
echo "bireal18 exit=$?"; head -c 1500 gpurun_out/bireal18.txt; tail -5 gpurun_out/bireal18.err
