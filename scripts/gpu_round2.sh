#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
echo "--- no-alloc test alone (fresh process)"
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_noalloc_alone.log
echo "--- full gpu suite"
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest_gpu.log
echo "--- ncu launch list"
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
echo "ncu exit=$?"; tail -3 gpurun_out/ncu.log; wc -l gpurun_out/launches.csv
