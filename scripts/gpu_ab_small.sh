#!/bin/bash
# small-batch A/B of an environment switch: bash scripts/gpu_ab_small.sh VAR=off_value
cd "$(dirname "$0")/.."
pr() { python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])"; }
for b in 128 512; do for g in "" "--graph"; do
echo -n "batch $b $g default: "; timeout 600 python bench.py --steps 50 --warmup 10 --batch $b --no-cpu-baseline $g 2>/dev/null | pr
echo -n "batch $b $g $1: "; env "$1" timeout 600 python bench.py --steps 50 --warmup 10 --batch $b --no-cpu-baseline $g 2>/dev/null | pr
done; done
