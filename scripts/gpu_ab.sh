#!/bin/bash
# parity tests + smoke, then the step with an environment switch on and off: bash scripts/gpu_ab.sh VAR=off_value
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -30 | tee gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5 | tee gpurun_out/smoke.log
pr() { python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'])"; }
for i in 1 2; do
timeout 900 python bench.py --steps 20 --warmup 5 2>/dev/null | { echo -n "default: "; pr; }
env "$1" timeout 900 python bench.py --steps 20 --warmup 5 2>/dev/null | { echo -n "$1: "; pr; }
done
timeout 900 python bench.py --steps 20 --warmup 5 --breakdown > gpurun_out/bench.json 2> gpurun_out/bench_breakdown.txt
grep TOTAL gpurun_out/bench_breakdown.txt | head -12
