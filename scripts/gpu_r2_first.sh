#!/bin/bash
# r2 first call: GPU tests (new full-size parity + convergence), smoke, bench, ncu of the HBM-bound kernels
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nproc > gpurun_out/nproc.txt; free -g | head -2 >> gpurun_out/nproc.txt
timeout 2400 python -m pytest tests -m gpu -q --durations=15 2>&1 | tail -60 | tee gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5 | tee gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 --breakdown > gpurun_out/bench.json 2> gpurun_out/bench_breakdown.txt
echo "bench exit=$?"; grep TOTAL gpurun_out/bench_breakdown.txt | head -12
timeout 1200 ncu --set full --clock-control none --import-source on \
  -k regex:"preprocess_kernel|bn_apply_sign_kernel|maxpool2_i16_kernel|bn_bwd_apply_kernel|bn_bwd_reduce_kernel|conv_first_bn_sign_kernel|conv_first_gram_kernel|adam_clip_kernel|bn_stats_kernel" \
  -s 40 -c 30 -o gpurun_out/prof_hbm python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_hbm.log 2>&1
echo "ncu exit=$?"; ls -la gpurun_out/*.ncu-rep
