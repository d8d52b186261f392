#!/bin/bash
# parity tests, smoke, bench with / without forward statistics in the GEMM epilogue
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -30 | tee gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5 | tee gpurun_out/smoke.log
pr() { python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'])"; }
for i in 1 2; do
timeout 900 python bench.py --steps 20 --warmup 5 2>/dev/null > gpurun_out/bench2.json; echo -n "fused fwd stats: "; pr < gpurun_out/bench2.json
ZKB_FUSE_FWD_STATS=0 timeout 900 python bench.py --steps 20 --warmup 5 2>/dev/null > gpurun_out/bench_nost.json; echo -n "separate bn_stats: "; pr < gpurun_out/bench_nost.json
done
timeout 900 python bench.py --steps 20 --warmup 5 --breakdown > gpurun_out/bench.json 2> gpurun_out/bench_breakdown.txt
grep TOTAL gpurun_out/bench_breakdown.txt | head -16
