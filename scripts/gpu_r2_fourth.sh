#!/bin/bash
# r2 fourth call (1 GPU): all GPU tests (incl. Bi-Real units), bench, preprocessing point + ncu
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q 2>&1 | tail -40 | tee gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 --breakdown > gpurun_out/bench.json 2> gpurun_out/bench_breakdown.txt
echo "bench exit=$?"; grep -E "maxpool|gram|preprocess" gpurun_out/bench_breakdown.txt | head -8
python -c "import json; d=json.load(open('gpurun_out/bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
timeout 300 python scripts/bench_preprocess.py > gpurun_out/preprocess_n1.json 2> gpurun_out/preprocess_n1.err; cat gpurun_out/preprocess_n1.json; tail -2 gpurun_out/preprocess_n1.err
timeout 600 ncu --set full --clock-control none -k regex:"preprocess_kernel" -s 6 -c 1 -o gpurun_out/prof_pre224 python scripts/bench_preprocess.py > gpurun_out/ncu_pre224.log 2>&1
echo "ncu2 exit=$?"
