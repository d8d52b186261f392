#!/bin/bash
# last seconds of the round's GPU budget: the BinaryNet main path after the per-channel kernels learnt
# row pitches / odd widths and the wgrad GEMM its 64-row units -- step parity, smoke, one short bench line
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 55 python -m pytest tests/test_gpu_train_step.py -m gpu -q -x -p no:cacheprovider \
  -k "train_step_matches_oracle or pool_in_forward_epilogue or fused_first_layer" 2>&1 | tail -6 | tee gpurun_out/last_tests.log
timeout 25 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/last_smoke.log
timeout 30 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-binary-gemm > gpurun_out/last_bench.json 2> gpurun_out/last_bench.err
echo "bench exit=$?"; python -c "import json; d=json.load(open('gpurun_out/last_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
