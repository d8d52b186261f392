#!/bin/bash
# r2 extras (1 GPU): validation-pass throughput, configs[0] batch on the GPU, CIFAR-shaped line of the final bench.py, preprocessing ncu
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python scripts/bench_eval.py 2>&1 | tail -3 | tee gpurun_out/eval_bench.txt
pr() { python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['config']['launch'])"; }
{
for b in 128 1024; do
echo -n "batch $b per-kernel: "; timeout 300 python bench.py --steps 50 --warmup 10 --batch $b --no-cpu-baseline --no-binary-gemm 2>/dev/null | pr
echo -n "batch $b graph:      "; timeout 300 python bench.py --steps 50 --warmup 10 --batch $b --no-cpu-baseline --no-binary-gemm --graph 2>/dev/null | pr
done
} | tee gpurun_out/small_batch.txt
timeout 300 python bench.py --config cifar --steps 50 --warmup 10 > gpurun_out/bench_cifar.json 2>/dev/null; python -c "import json; d=json.load(open('gpurun_out/bench_cifar.json')); print('cifar', d['value'], d['ms_per_step'], d['config'])"
timeout 300 python scripts/bench_preprocess.py > gpurun_out/preprocess_n1.json 2>/dev/null; cat gpurun_out/preprocess_n1.json
timeout 300 ncu --set full --clock-control none -k regex:"preprocess_kernel" -s 6 -c 1 -o gpurun_out/prof_pre224_final python scripts/bench_preprocess.py > gpurun_out/ncu_pre224.log 2>&1
echo "ncu exit=$?"
