#!/bin/bash
# launch list (shares) + one full ncu capture of the two dominant GEMMs (layer 1's fused dgrad and wgrad)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
echo "launch list exit=$?"
python scripts/summarize_launches.py gpurun_out/launches.csv --last-step > gpurun_out/launches_summary.txt; head -30 gpurun_out/launches_summary.txt
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 94 -c 2 -o gpurun_out/prof_bwd1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit=$?"
