"""Train-step throughput of the two larq-zoo models BASELINE.json names, at its shapes: configs[3]
BinaryAlexNet and configs[4] Bi-Real-Net-18, 224x224x3 images, 1000 classes, 256 images per GPU
(SURVEY.md section 8d "C4" / "C5"; both models are larq-zoo's, [EXT], restated in
oracle/alexnet_oracle.py and oracle/bireal_oracle.py).  One GPU; synthetic uint8 pool, random-init weights.

  python scripts/bench_zoo.py --model bireal18|alexnet [--batch 256] [--steps 10] [--warmup 3]
                              [--breakdown] [--cpu-baseline]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
         scripts/bench_zoo.py --model ...        # data-parallel: --batch per GPU (weak scaling), gradients
                                                 # summed with NCCL, BatchNorm statistics per replica

A step = fused preprocessing (zkb_preprocess_u8: gather, random crop 256 -> 224, flip, u8 -> f32)
-> forward -> loss -> full backward -> Adam (+ latent-weight clip).  Timed with CUDA events on the
launching stream after the warm-up; the activations of one step (several GB) exceed the 126 MB L2,
no flush.  --breakdown adds one pass with an event pair around every launch (serialised; shares, not
absolutes).  --cpu-baseline times the float32 torch-CPU oracle of the same model on a bounded
sample (batch 16, 2 steps, all host threads): a reported baseline, not a target."""
import argparse
import json
import os
import sys
import time
from collections import defaultdict

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from zookeeper_b200 import native as zn  # noqa: E402
from zookeeper_b200.alexnet import BinaryAlexNetEngine  # noqa: E402
from zookeeper_b200.bireal import BiRealNetEngine, bireal18_plan  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--image", type=int, default=224)
ap.add_argument("--breakdown", action="store_true")
ap.add_argument("--model", choices=("bireal18", "alexnet"), default="bireal18")
ap.add_argument("--cpu-baseline", action="store_true")
args = ap.parse_args()

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
if world > 1:
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
B, O, S, NC = args.batch, args.image, args.image + 32, 1000
rng = np.random.default_rng(0)          # the same initial weights on every replica


def bireal_params(nc):
    # random-init weights (Glorot-uniform kernels, BatchNorm at identity), written through the public loader
    p, c = {}, 64
    lim = np.sqrt(6.0 / (49 * (3 + c)))
    p["stem_kernel"] = rng.uniform(-lim, lim, (7, 7, 3, c)).astype(np.float32)
    p["stem_gamma"], p["stem_beta"] = np.ones(c, np.float32), np.zeros(c, np.float32)
    for u, kind in enumerate(bireal18_plan()):
        co = 2 * c if kind == "down" else c
        lim = np.sqrt(6.0 / (9 * (c + co)))
        p[f"kernel{u}"] = rng.uniform(-lim, lim, (3, 3, c, co)).astype(np.float32)
        p[f"gamma{u}"], p[f"beta{u}"] = np.ones(co, np.float32), np.zeros(co, np.float32)
        if kind == "down":
            lim = np.sqrt(6.0 / (c + co))
            p[f"sc_kernel{u}"] = rng.uniform(-lim, lim, (c, co)).astype(np.float32)
            p[f"sc_gamma{u}"], p[f"sc_beta{u}"] = np.ones(co, np.float32), np.zeros(co, np.float32)
        c = co
    lim = np.sqrt(6.0 / (c + nc))
    p["dense_kernel"] = rng.uniform(-lim, lim, (c, nc)).astype(np.float32)
    p["dense_bias"] = np.zeros(nc, np.float32)
    return p


if args.model == "bireal18":
    eng = BiRealNetEngine(bireal18_plan(), (O, O), B, NC)
    params = bireal_params(NC)
    eng.load_params(params)
    call_host = eng.body
    # algorithmic work per image (SURVEY.md section 8d conventions: binary 2*MAC ops, padded taps count)
    hw, c, bin_mac = (O // 4) ** 2, 64, 0
    for kind in bireal18_plan():
        if kind == "down":
            hw, co = hw // 4, 2 * c
        else:
            co = c
        bin_mac += hw * 9 * c * co
        c = co
    float_mac = (O // 2) ** 2 * 147 * 64
    name = "Bi-Real-Net-18"
else:
    eng = BinaryAlexNetEngine((O, O), B, NC)
    params = {}
    for L in eng.L:
        shape = eng._kshape(L)
        fan = (L.k * L.k * (L.cin + L.cout)) if L.kind == "conv" else (L.cin + L.cout)
        lim = np.sqrt(6.0 / fan)
        params[f"kernel{L.idx}"] = rng.uniform(-lim, lim, shape).astype(np.float32)
        params[f"beta{L.idx}"] = np.zeros(L.cout, np.float32)
    eng.load_params(params)
    call_host = eng
    bin_mac = sum((L.conv_hw[0] * L.conv_hw[1] * L.k * L.k if L.kind == "conv" else 1) * L.cin * L.cout for L in eng.L[1:])
    L0 = eng.L[0]
    float_mac = L0.conv_hw[0] * L0.conv_hw[1] * L0.k * L0.k * 3 * L0.cout
    name = "BinaryAlexNet"

n_pool = 2 * B
g = torch.Generator(device="cpu").manual_seed(1234 + rank)
pool = torch.randint(0, 256, (n_pool, S, S, 3), generator=g, dtype=torch.uint8).to(dev)
labels = torch.randint(0, NC, (n_pool,), generator=g, dtype=torch.int64).to(dev)
index = [torch.randperm(n_pool, generator=g)[:B].to(torch.int32).to(dev) for _ in range(4)]
lab_out = torch.empty(B, dtype=torch.int32, device=dev)
stream = torch.cuda.current_stream().cuda_stream


def step(k):
    zn.call("zkb_preprocess_u8", zn.ptr(pool), zn.ptr(labels), zn.ptr(index[k % 4]), 0, zn.ptr(eng.x), zn.ptr(lab_out),
            0, B, S, S, 3, O, O, S, 1, 0, NC, 1234 + rank, k, 0, stream)
    eng.train_step(eng.x, lab_out)


for k in range(args.warmup):
    step(k)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for k in range(args.steps):
    step(args.warmup + k)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / args.steps
loss = eng.last_loss()
in_sync = None
if world > 1:
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    # every replica applied the same summed gradients to the same initial weights: the arenas must agree
    arenas = [eng.params] if args.model == "alexnet" else [eng.body.params, eng.params]
    chk = torch.stack([a.double().sum() for a in arenas] + [a.double().abs().sum() for a in arenas])
    allc = [torch.zeros_like(chk) for _ in range(world)]
    dist.all_gather(allc, chk)
    in_sync = all(bool(torch.equal(allc[0], c)) for c in allc)

line = {"metric": f"{name} train images/sec", "value": world * B / (ms * 1e-3), "unit": "images/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "dtype": "int8 / bf16 +-1 fwd (exact) / bf16x2 bwd (f32 acc) / f32 stem, shortcuts and elementwise", "data": "synthetic",
        "config": {"workload": f"{name} (larq-zoo, [EXT]) {O}x{O}x3 synthetic, {NC} classes, batch {B}/GPU",
                   "global_batch": world * B, "parallelism": f"dp{world}", "batchnorm": "per-replica",
                   "launch": "per-kernel launches"},
        "scaling": "weak", "replicas_in_sync": in_sync,
        "gpu_launches": (eng.launches_per_step + 1) * args.steps, "last_loss": loss,
        "work_per_image": {"binary_mac": bin_mac, "stem_float_mac": float_mac},
        "binary_fwd_pops_if_all_time": 2 * bin_mac * world * B / (ms * 1e-3) / 1e15}
if args.cpu_baseline and rank == 0:
    nt = os.cpu_count() or 1
    torch.set_num_threads(nt)
    cb, csteps = 16, 2
    crng = np.random.default_rng(1)
    cimg = crng.uniform(-1, 1, (cb, O, O, 3)).astype(np.float32)
    clab = crng.integers(0, NC, (cb,))
    if args.model == "bireal18":
        from oracle import bireal_oracle as BO
        run = lambda: BO.forward_backward(params, bireal18_plan(), cimg, clab, torch.float32)   # noqa: E731
    else:
        from oracle import alexnet_oracle as AO
        run = lambda: AO.forward_backward(params, cimg, clab, torch.float32)                     # noqa: E731
    run()
    t0 = time.perf_counter()
    for _ in range(csteps):
        run()
    dt = (time.perf_counter() - t0) / csteps
    line["cpu_baseline"] = {"value": cb / dt, "unit": "images/s", "cores": nt, "kind": "port",
                            "sample": f"{csteps} forward+backward passes of batch {cb} of the torch-CPU fp32 restatement (no optimizer step), {nt} threads"}
if rank == 0:
    print(json.dumps(line), flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()

if args.breakdown and world == 1:
    per = defaultdict(list)
    orig = call_host._call

    def timed(name, *a):
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        orig(name, *a)
        a1.record()
        per[name].append((a0, a1))

    call_host._call = timed
    step(0)
    torch.cuda.synchronize()
    tot = {k: sum(x.elapsed_time(y) for x, y in v) for k, v in per.items()}
    allms = sum(tot.values())
    print(f"per-launch CUDA-event pass: {allms:.3f} ms in {sum(len(v) for v in per.values())} launches")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        print(f"  {k:32s} {v:8.3f} ms {100 * v / allms:5.1f}%  x{len(per[k])}")
