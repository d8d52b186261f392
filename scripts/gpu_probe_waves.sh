#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
pr() { python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])"; }
{
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q 2>&1 | tail -2
echo "=== wgrad alone"; timeout 300 python scripts/bench_gemm.py wgrad
for w in 2 1 3 2 1; do
echo -n "step, wgrad waves=$w: "; ZKB_WGRAD_WAVES=$w timeout 600 python bench.py --steps 20 --warmup 5 2>/dev/null | pr
done
} 2>&1 | tee gpurun_out/probe_waves.log
