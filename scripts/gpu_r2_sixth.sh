#!/bin/bash
# r2 sixth call (1 GPU): GPU tests with the fused conv+pool forward, bench A/B of the fusion
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q 2>&1 | tail -30 | tee gpurun_out/pytest_gpu.log
pr() { python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'])"; }
for i in 1 2; do
echo -n "fused pool: "; timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-binary-gemm 2>/dev/null | pr
echo -n "separate pool: "; ZKB_FUSE_POOL=0 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-binary-gemm 2>/dev/null | pr
done
timeout 900 python bench.py --steps 20 --warmup 5 --breakdown --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench_breakdown.txt
grep -E "fwd1|fwd5|fwd3" gpurun_out/bench_breakdown.txt | head -12
