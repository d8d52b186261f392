#!/bin/bash
# r2 widening (1 GPU): narrow-channel / 5x5 binary convs, float stems, Bi-Real-Net-18 and BinaryAlexNet steps; then their bench lines
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_alexnet.py tests/test_gpu_stem_and_narrow.py tests/test_gpu_bireal.py -m gpu -q -p no:cacheprovider \
  --timeout 240 -o timeout_method=thread -rf 2>&1 | tail -80 | tee gpurun_out/widen_tests.log
for m in bireal18 alexnet; do
timeout 400 python scripts/bench_zoo.py --model $m --breakdown --cpu-baseline > gpurun_out/zoo_$m.txt 2> gpurun_out/zoo_$m.err
echo "$m exit=$?"; head -c 1800 gpurun_out/zoo_$m.txt; tail -5 gpurun_out/zoo_$m.err
done
