#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_train_step.py -m gpu -x -q 2>&1 | tail -40 | tee gpurun_out/pytest_gpu.log
