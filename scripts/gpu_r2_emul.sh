#!/bin/bash
# Ceiling of a half-cost second GEMM term (e4m3 residual on kind::f8f6f4 at twice the bf16 rate): a
# -DZKB_MEASURE build issues half of the second term's MMAs (ZKB_GEMM_DBG bit 4) and, in swap mode,
# loads half of its bytes (bit 8).  Timing only -- the results of such launches are wrong.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
ZKB_MEASURE=1 python -m zookeeper_b200.build --force | tail -1
echo "=== baseline (measure build, no switch)"; timeout 300 python scripts/bench_gemm.py dgrad wgrad
echo "=== second term: half the MMAs (ZKB_GEMM_DBG=4)"; ZKB_GEMM_DBG=4 timeout 300 python scripts/bench_gemm.py dgrad wgrad
echo "=== second term: half the MMAs and, in swap mode, half the bytes (ZKB_GEMM_DBG=12)"; ZKB_GEMM_DBG=12 timeout 300 python scripts/bench_gemm.py dgrad wgrad
pr() { python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])"; }
echo -n "=== step, baseline: "; timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-binary-gemm 2>/dev/null | pr
echo -n "=== step, ZKB_GEMM_DBG=12 (timing only): "; ZKB_GEMM_DBG=12 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-binary-gemm 2>/dev/null | pr
python -m zookeeper_b200.build --force | tail -1
} 2>&1 | tee gpurun_out/emul_half_cost_second_term.txt
