#!/bin/bash
# r2 seventh call (1 GPU): all GPU tests, CIFAR-shaped bench after the first-layer kernel rewrite
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q 2>&1 | tail -15 | tee gpurun_out/pytest_gpu.log
timeout 600 python bench.py --config cifar --steps 50 --warmup 10 --breakdown > gpurun_out/bench_cifar.json 2> gpurun_out/bench_cifar_breakdown.txt
echo "cifar exit=$?"; head -12 gpurun_out/bench_cifar_breakdown.txt
python -c "import json; d=json.load(open('gpurun_out/bench_cifar.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
timeout 600 python bench.py --config cifar --steps 50 --warmup 10 --graph --no-cpu-baseline --no-binary-gemm > gpurun_out/bench_cifar_graph.json 2>/dev/null
python -c "import json; d=json.load(open('gpurun_out/bench_cifar_graph.json')); print('graph', d['value'], d['ms_per_step'])"
