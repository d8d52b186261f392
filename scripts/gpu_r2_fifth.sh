#!/bin/bash
# r2 fifth call (1 GPU): all GPU tests, preprocessing point, small-batch per-kernel vs graph (tensor-map cache)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q 2>&1 | tail -40 | tee gpurun_out/pytest_gpu.log
timeout 300 python scripts/bench_preprocess.py > gpurun_out/preprocess_n1.json 2> gpurun_out/preprocess_n1.err; cat gpurun_out/preprocess_n1.json; tail -2 gpurun_out/preprocess_n1.err
timeout 600 python bench.py --batch 128 --steps 50 --warmup 10 --no-cpu-baseline --no-binary-gemm 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('b128 per-kernel', d['value'], d['ms_per_step'])"
timeout 600 python bench.py --batch 128 --steps 50 --warmup 10 --no-cpu-baseline --no-binary-gemm --graph 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('b128 graph', d['value'], d['ms_per_step'])"
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-binary-gemm 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('b4096', d['value'], d['ms_per_step'], d['e2e']['value'])"
