#!/bin/bash
# final 8-GPU call: weak bench line of the final code, CIFAR-shaped config (256 per GPU), preprocessing replicas at 4 and 8
N=${1:-8}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node $N --master-port 29561 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_weak_n$N.json 2> gpurun_out/bench_weak_n$N.err
echo "weak exit=$?"; python -c "import json; d=json.load(open('gpurun_out/bench_weak_n$N.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['allreduce']['exposed_ms_per_step'])"
timeout 600 $TR --nproc-per-node $N --master-port 29562 bench.py --gpus $N --config cifar --steps 50 --warmup 10 > gpurun_out/bench_cifar_n$N.json 2> gpurun_out/bench_cifar_n$N.err
echo "cifar exit=$?"; python -c "import json; d=json.load(open('gpurun_out/bench_cifar_n$N.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['allreduce']['exposed_ms_per_step'])"
timeout 600 $TR --nproc-per-node $N --master-port 29563 bench.py --gpus $N --config cifar --graph --steps 50 --warmup 10 > gpurun_out/bench_cifar_graph_n$N.json 2> gpurun_out/bench_cifar_graph_n$N.err
echo "cifar graph exit=$?"; python -c "import json; d=json.load(open('gpurun_out/bench_cifar_graph_n$N.json')); print(d['value'], d['ms_per_step'])"
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-binary-gemm 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('n1 on this box', d['value'], d['ms_per_step'])" | tee gpurun_out/bench_n1_same_box.txt
timeout 300 $TR --nproc-per-node 4 --master-port 29564 scripts/bench_preprocess.py > gpurun_out/preprocess_n4.json 2> gpurun_out/preprocess_n4.err; cat gpurun_out/preprocess_n4.json
timeout 300 $TR --nproc-per-node $N --master-port 29565 scripts/bench_preprocess.py > gpurun_out/preprocess_n$N.json 2> gpurun_out/preprocess_n$N.err; cat gpurun_out/preprocess_n$N.json
tail -2 gpurun_out/bench_cifar_graph_n$N.err
