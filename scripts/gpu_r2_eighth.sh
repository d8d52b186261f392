#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_bireal.py -m gpu -q 2>&1 | tail -25 | tee gpurun_out/pytest_bireal.log
