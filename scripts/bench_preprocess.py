"""Fused-Preprocessing throughput at an ImageNet-shaped input (BASELINE.json configs[4]: 224x224x3,
256 images per GPU, "fused Preprocessing throughput sweep at 1/2/4/8 B200").

  python scripts/bench_preprocess.py                                   # one GPU
  python -m torch.distributed.run --nproc-per-node N ... scripts/bench_preprocess.py   # N replicas

A pass = zkb_preprocess_u8 over one batch: gather by index from a resident uint8 pool of 256x256x3
images, random 224x224 crop, random flip, u8 -> f32 (x / 127.5 - 1), labels.  The path shards by
sample with no exchange step, so N GPUs are N independent replicas; the time is the max over ranks.
Algorithmic bytes per image: 224*224*3 * (1 B read + 4 B written) + 8 B label
= 752 648 B.  Working set per pass (50 MB pool slice read sparsely + 154 MB written, rotating over
4 pool batches and 2 output buffers) exceeds the 126 MB L2; no flush."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from zookeeper_b200 import native as zn  # noqa: E402

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)

B, S, O, C = int(os.environ.get("B", "256")), 256, 224, 3
n_pool = 4 * B
g = torch.Generator(device="cpu").manual_seed(1234 + rank)
pool = torch.randint(0, 256, (n_pool, S, S, C), generator=g, dtype=torch.uint8).to(dev)
labels = torch.randint(0, 1000, (n_pool,), generator=g, dtype=torch.int64).to(dev)
index = [torch.randperm(n_pool, generator=g)[:B].to(torch.int32).to(dev) for _ in range(8)]
outs = [torch.empty(B, O, O, C, dtype=torch.float32, device=dev) for _ in range(2)]
lab_out = torch.empty(B, dtype=torch.int32, device=dev)
stream = torch.cuda.current_stream().cuda_stream


def one(k):
    zn.call("zkb_preprocess_u8", zn.ptr(pool), zn.ptr(labels), zn.ptr(index[k % 8]), 0, zn.ptr(outs[k % 2]),
            zn.ptr(lab_out), 0, B, S, S, C, O, O, S, 1, 0, 1000, 1234 + rank, k, 0, stream)


for k in range(5):
    one(k)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
n = 50
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for k in range(n):
    one(k)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
if world > 1:
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
bytes_per_img = O * O * C * 5 + 8
peak = 6555.5
try:
    with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
        peak = json.load(f).get("hbm_gbs", peak)
except OSError:
    pass
if rank == 0:
    gbs = bytes_per_img * B / (ms * 1e-3) / 1e9
    print(json.dumps({"metric": "fused Preprocessing throughput", "workload": f"{S}x{S}x{C} u8 pool -> {O}x{O}x{C} f32, batch {B}/GPU",
                      "n_gpus": world, "ms_per_pass": ms, "images_per_s": world * B / (ms * 1e-3),
                      "gbs_per_gpu": gbs, "gbs_total": gbs * world, "hbm_peak_gbs": peak, "frac": gbs / peak,
                      "bytes_per_image": bytes_per_img, "scaling": "weak (independent replicas, no collective)"}))
if world > 1:
    dist.destroy_process_group()
