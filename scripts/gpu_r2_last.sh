#!/bin/bash
# the BinaryNet main path after the per-channel kernels learnt row pitches / odd widths and the wgrad
# GEMM its 64-row units: kernel-level tests, step parity, the allocation-free entry point, smoke, one
# short bench line (run in three calls at the end of round 2, when little GPU budget was left)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_zz_no_alloc.py tests/test_gpu_train_step.py -m gpu -q -p no:cacheprovider \
  --deselect tests/test_gpu_train_step.py::test_training_converges_like_the_oracle_and_checkpoint_roundtrip 2>&1 | tail -6 | tee gpurun_out/last_tests.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/last_smoke.log
timeout 60 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-binary-gemm > gpurun_out/last_bench.json 2> gpurun_out/last_bench.err
echo "bench exit=$?"; python -c "import json; d=json.load(open('gpurun_out/last_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
