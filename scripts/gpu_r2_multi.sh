#!/bin/bash
# multi-GPU call: N = $1 GPUs of one box -- SyncBN parity (test + kept log), weak / strong / SyncBN bench lines, preprocessing replicas
N=${1:-2}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_multi.py -q -m gpu 2>&1 | tail -5 | tee gpurun_out/pytest_multi_n$N.log
timeout 600 $TR --master-port 29551 scripts/dp_parity.py 2>&1 | grep -v "^W\|^\[W\|Warning" | tail -8 | tee gpurun_out/dp_parity_n$N.stdout
timeout 600 $TR --master-port 29552 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_weak_n$N.json 2> gpurun_out/bench_weak_n$N.err
echo "weak exit=$?"; python -c "import json; d=json.load(open('gpurun_out/bench_weak_n$N.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d.get('allreduce'))"
timeout 600 $TR --master-port 29553 bench.py --gpus $N --scaling strong --steps 30 --warmup 5 > gpurun_out/bench_strong_n$N.json 2> gpurun_out/bench_strong_n$N.err
echo "strong exit=$?"; python -c "import json; d=json.load(open('gpurun_out/bench_strong_n$N.json')); print(d['value'], d['ms_per_step'], d['config'], d.get('allreduce'))"
timeout 600 $TR --master-port 29554 bench.py --gpus $N --sync-bn --steps 20 --warmup 5 > gpurun_out/bench_syncbn_n$N.json 2> gpurun_out/bench_syncbn_n$N.err
echo "syncbn exit=$?"; python -c "import json; d=json.load(open('gpurun_out/bench_syncbn_n$N.json')); print(d['value'], d['ms_per_step'], d['config']['batchnorm'])"
timeout 300 $TR --master-port 29555 scripts/bench_preprocess.py > gpurun_out/preprocess_n$N.json 2> gpurun_out/preprocess_n$N.err; cat gpurun_out/preprocess_n$N.json
tail -3 gpurun_out/bench_weak_n$N.err
