#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
echo "--- full gpu suite"
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 | tee gpurun_out/pytest_gpu.log
echo "--- example on GPU (1 epoch, synthetic)"
timeout 600 python examples/larq_experiment.py BinaryNetMnist epochs=2 batch_size=1024 dataset.num_train_examples=8192 dataset.num_validation_examples=2048 2>&1 | tail -6 | tee gpurun_out/example.log
echo "--- ncu full on wgemm + kgemm"
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 28 -c 6 -o gpurun_out/prof_r1_gemm python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu exit=$?"; tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/
