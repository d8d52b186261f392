#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
for b in 128 512 1024; do
echo "== batch=$b"; timeout 300 python bench.py --steps 50 --warmup 10 --no-cpu-baseline --batch $b | cut -c1-170
done
echo "== example default (batch 128), 1 epoch of 12800 samples"
timeout 300 python examples/larq_experiment.py BinaryNetMnist epochs=2 dataset.num_train_examples=12800 dataset.num_validation_examples=1280 2>&1 | tail -2
} 2>&1 | tee gpurun_out/ab.log
