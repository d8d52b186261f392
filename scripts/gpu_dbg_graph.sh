#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
for i in 1 2 3; do
echo "--- fused stats on, run $i"; timeout 300 python -m pytest tests/test_gpu_train_step.py -x -q -k graph_replay 2>&1 | grep -E "passed|failed|assert 0\."
done
for i in 1 2 3; do
echo "--- fused stats off, run $i"; ZKB_FUSE_FWD_STATS=0 timeout 300 python -m pytest tests/test_gpu_train_step.py -x -q -k graph_replay 2>&1 | grep -E "passed|failed|assert 0\."
done
timeout 300 python - <<'PY'
import numpy as np, torch, sys
sys.path.insert(0, ".")
from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers
B, n = 64, 512
rng = np.random.default_rng(0)
pool = torch.as_tensor(rng.integers(0, 256, (n, 28, 28, 1), dtype=np.uint8)).cuda()
labs = torch.as_tensor(rng.integers(0, 10, (n,)).astype(np.int64)).cuda()
idx = [torch.as_tensor(rng.permutation(n)[:B].astype(np.int32)).cuda() for _ in range(4)]
def run(graph):
    e = BinaryNetEngine(binarynet_layers(), (28, 28, 1), B, 10)
    e.init_glorot(0)
    out = []
    if graph:
        e.train_step(pool, labs, idx[0], pad_size=32, seed=5); out.append(e.last_loss())
        e.capture_train_step(pool, labs, 32, 5)
        for k in range(1, 4):
            e.train_step_graph(idx[k]); out.append(e.last_loss())
    else:
        for k in range(4):
            e.train_step(pool, labs, idx[k], pad_size=32, seed=5); out.append(e.last_loss())
    return out
for g in (False, False, True, True):
    print("graph" if g else "plain", ["%.6f" % v for v in run(g)])
PY
} 2>&1 | tee gpurun_out/dbg_graph.log
