#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
timeout 900 python -m pytest tests/test_gpu_train_step.py -m gpu -x -q 2>&1 | tail -12
for b in 128 1024 4096; do
echo "== batch=$b eager"; timeout 300 python bench.py --steps 50 --warmup 10 --no-cpu-baseline --batch $b | cut -c1-170
echo "== batch=$b graph"; timeout 300 python bench.py --steps 50 --warmup 10 --no-cpu-baseline --batch $b --graph | cut -c1-170
done
echo "== example default (batch 128)"
timeout 300 python examples/larq_experiment.py BinaryNetMnist epochs=2 dataset.num_train_examples=12800 dataset.num_validation_examples=1280 2>&1 | tail -2
} 2>&1 | tee gpurun_out/ab.log
