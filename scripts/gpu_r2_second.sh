#!/bin/bash
# r2 second call: GPU tests, bench line with the new keys, CIFAR-shaped config
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q --durations=8 2>&1 | tail -40 | tee gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 --breakdown > gpurun_out/bench.json 2> gpurun_out/bench_breakdown.txt
echo "bench exit=$?"; tail -c 3000 gpurun_out/bench.json
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
echo "ref exit=$?"; cat gpurun_out/bench_ref.json
timeout 600 python bench.py --config cifar --steps 50 --warmup 10 --breakdown > gpurun_out/bench_cifar.json 2> gpurun_out/bench_cifar_breakdown.txt
echo "cifar exit=$?"; cut -c1-400 gpurun_out/bench_cifar.json
timeout 600 python bench.py --config cifar --steps 50 --warmup 10 --graph --no-cpu-baseline --no-binary-gemm > gpurun_out/bench_cifar_graph.json 2>/dev/null
echo "cifar graph exit=$?"; cut -c1-300 gpurun_out/bench_cifar_graph.json
