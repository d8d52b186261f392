#!/bin/bash
# One GPU visit: parity tests, smoke, bench with the first-layer backward fused and separate.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -30 | tee gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5 | tee gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 --breakdown > gpurun_out/bench.json 2> gpurun_out/bench_breakdown.txt
echo "bench exit=$?"; cat gpurun_out/bench.json; grep TOTAL gpurun_out/bench_breakdown.txt | head -14
ZKB_FUSE_FIRST_BWD=0 timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_unfused.json 2>/dev/null
echo "unfused:"; cat gpurun_out/bench_unfused.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench2.json 2>/dev/null
echo "fused again:"; cat gpurun_out/bench2.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])"
