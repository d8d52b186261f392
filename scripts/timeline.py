"""Per-launch start/end times of one overlapped train step (both streams), batch 4096."""
import sys

import torch

sys.path.insert(0, ".")
from zookeeper_b200 import native as zn  # noqa: E402
from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers  # noqa: E402

B = 4096
eng = BinaryNetEngine(binarynet_layers(), (28, 28, 1), B, 10)
eng.init_glorot(0)
img = torch.randint(0, 256, (B, 28, 28, 1), dtype=torch.uint8, device="cuda")
lab = torch.randint(0, 10, (B,), dtype=torch.int64, device="cuda")
for _ in range(5):
    eng.train_step(img, lab, None, pad_size=32, seed=1)
torch.cuda.synchronize()
eng.watch = set(zn.SIGNATURES)          # a set keeps the two-stream schedule
eng.watched = []
t0 = torch.cuda.Event(enable_timing=True)
t0.record()
eng.train_step(img, lab, None, pad_size=32, seed=1)
torch.cuda.synchronize()
rows = []
for name, tag, e0, e1 in eng.watched:
    rows.append((t0.elapsed_time(e0), t0.elapsed_time(e1), name.replace("zkb_", ""), tag))
rows.sort()
end = max(r[1] for r in rows)
print(f"step end at {end:.3f} ms")
for s, e, name, tag in rows:
    side = "SIDE" if "wgrad" in name and "first" not in name else "main"
    print(f"{s:8.3f} {e:8.3f} {e - s:7.3f}  {side}  {name:24s} {tag}")
