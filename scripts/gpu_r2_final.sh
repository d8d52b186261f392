#!/bin/bash
# r2 final call (1 GPU): full GPU test suite, smoke, the bench line, the ncu launch list of the same
# command and one full capture of the two dominant backward GEMMs (layer 1: fused dgrad, wgrad)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q 2>&1 | tail -12 | tee gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | tee gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 --breakdown > gpurun_out/bench.json 2> gpurun_out/bench_breakdown.txt
echo "bench exit=$?"; python -c "import json; d=json.load(open('gpurun_out/bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['binary_gemm']['frac'])"
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-binary-gemm > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-binary-gemm > gpurun_out/ncu.log 2>&1
echo "launch list exit=$?"
python scripts/summarize_launches.py gpurun_out/launches.csv --last-step > gpurun_out/launches_summary.txt; head -14 gpurun_out/launches_summary.txt
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"kgemm_first_bf16|wgemm_bf16" -s 34 -c 2 -o gpurun_out/prof_bwd1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-binary-gemm > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit=$?"
timeout 600 ncu --set full --clock-control none -k regex:"kgemm_bf16_kernel" -s 30 -c 10 -o gpurun_out/prof_fwd python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-binary-gemm > gpurun_out/ncu_fwd.log 2>&1
echo "ncu fwd exit=$?"
