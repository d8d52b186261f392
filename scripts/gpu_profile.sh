#!/bin/bash
# launch list of one step (shares); the full capture of the dominant kernels is scripts/gpu_profile2.sh
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
echo "launch list exit=$?"
python scripts/summarize_launches.py gpurun_out/launches.csv --last-step > gpurun_out/launches_summary.txt; head -30 gpurun_out/launches_summary.txt
