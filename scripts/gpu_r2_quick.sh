#!/bin/bash
# quick 1-GPU check of the test files named on the command line (default: the zoo-model tests)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
T=${@:-tests/test_gpu_alexnet.py}
timeout 900 python -m pytest $T -m gpu -q -p no:cacheprovider --timeout 240 -o timeout_method=thread -rf 2>&1 | tail -120 | tee gpurun_out/quick_tests.log
