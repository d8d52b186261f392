#!/bin/bash
# r2 third call (1 GPU): kernel tests after the elementwise rewrites, bench, preprocessing sweep point, ncu of the rewritten kernels
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x 2>&1 | tail -30 | tee gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 --breakdown > gpurun_out/bench.json 2> gpurun_out/bench_breakdown.txt
echo "bench exit=$?"; grep -E "maxpool|gram|preprocess|bn_bwd_reduce|bn_apply_sign  " gpurun_out/bench_breakdown.txt | head -20
python -c "import json; d=json.load(open('gpurun_out/bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
timeout 300 python scripts/bench_preprocess.py > gpurun_out/preprocess_n1.json 2> gpurun_out/preprocess_n1.err; cat gpurun_out/preprocess_n1.json; tail -2 gpurun_out/preprocess_n1.err
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:"preprocess_kernel|maxpool2_i16_kernel|conv_first_gram_kernel|bn_bwd_reduce_kernel" \
  -s 12 -c 12 -o gpurun_out/prof_hbm2 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-binary-gemm > gpurun_out/ncu_hbm2.log 2>&1
echo "ncu exit=$?"
timeout 600 ncu --set full --clock-control none -k regex:"preprocess_kernel" -s 6 -c 1 -o gpurun_out/prof_pre224 python scripts/bench_preprocess.py > gpurun_out/ncu_pre224.log 2>&1
echo "ncu2 exit=$?"
