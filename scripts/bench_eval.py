"""Validation-pass throughput (images/s) of eval_step at batch 4096: folded thresholds vs BatchNorm passes."""
import os
import sys

import torch

sys.path.insert(0, ".")
from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers  # noqa: E402

B = int(os.environ.get("B", "4096"))
img = torch.randint(0, 256, (B, 28, 28, 1), dtype=torch.uint8, device="cuda")
lab = torch.randint(0, 10, (B,), dtype=torch.int64, device="cuda")
for fold in ("1", "0"):
    os.environ["ZKB_FOLD_EVAL"] = fold
    eng = BinaryNetEngine(binarynet_layers(), (28, 28, 1), B, 10)
    eng.init_glorot(0)
    for _ in range(3):
        eng.eval_step(img, lab)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    n = 20
    for _ in range(n):
        eng.eval_step(img, lab)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print(f"fold_eval={fold}: {ms:.3f} ms/step  {B / ms * 1e3:,.0f} images/s  ({eng.launches_per_step or eng._n_launch} launches)")
    del eng
