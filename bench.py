#!/usr/bin/env python
"""BinaryNet train-step benchmark (BASELINE.json metric: train images/sec; configs[1]).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # CPU restatement of the TF/Larq path

One JSON line on stdout (rank 0).  A "step" is one full train step of the example's BinaryNet on
one batch of synthetic MNIST-shaped data: fused preprocessing -> forward -> softmax-CE -> backward
-> (NCCL all-reduce) -> Adam + latent clip.  `value` times steps whose uint8 inputs are already
resident in HBM; `e2e` times the same steps fed from pinned HOST buffers, with the host->device copy
of every batch and the device->host read of the loss inside the timed region.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "BinaryNet train images/sec"
UNIT = "images/s"
INPUT_SHAPE = (28, 28, 1)
NUM_CLASSES = 10
PAD_SIZE = 32
BATCH_PER_GPU = 4096      # BASELINE.json configs[1]
CPU_SAMPLE_BATCH = 128    # BASELINE.json configs[0] (the reference's own CPU-runnable case)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=BATCH_PER_GPU, help="images per GPU per step")
    ap.add_argument("--breakdown", action="store_true", help="also print a per-kernel time table (stderr)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sync-bn", action="store_true")
    ap.add_argument("--graph", action="store_true", help="replay a captured CUDA graph per step (small batches)")
    return ap.parse_args()


def workload_name(batch):
    return f"BinaryNet (examples/larq_experiment.py) MNIST-shaped 28x28x1 synthetic, batch {batch}/GPU"


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle restatement of the reference's TF/Larq path, timed on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_train_steps(n_steps, warmup, batch, budget_s=None):
    """Time full train steps (preprocess + fwd + bwd + Adam + clip, fp32) of the CPU restatement."""
    import numpy as np
    import torch

    from oracle import binarynet_oracle as O

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    layers = O.binarynet_layers(INPUT_SHAPE, 128, 1024, NUM_CLASSES)
    p = O.glorot_uniform_params(layers, 0)
    m = {k: np.zeros_like(v) for k, v in p.items()}
    v = {k: np.zeros_like(v_) for k, v_ in p.items()}
    rng = np.random.default_rng(1234)
    img = rng.integers(0, 256, (batch,) + INPUT_SHAPE, dtype=np.uint8)
    lab = rng.integers(0, NUM_CLASSES, (batch,)).astype(np.int64)
    times = []
    t_begin = time.perf_counter()
    for s in range(warmup + n_steps):
        t0 = time.perf_counter()
        aug = O.augmentation_params(1234, s, batch, PAD_SIZE, INPUT_SHAPE[0], INPUT_SHAPE[1])
        x = O.preprocess(img, INPUT_SHAPE[:2], True, PAD_SIZE, aug)
        r = O.forward_backward(p, layers, x, lab, torch.float32)
        p, m, v = O.adam_clip_step(p, r.grads, m, v, s + 1, dtype=np.float32)
        dt = time.perf_counter() - t0
        if s >= warmup:
            times.append(dt)
        if budget_s is not None and s >= warmup and time.perf_counter() - t_begin > budget_s:
            break
    return times, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    times, cores = cpu_train_steps(args.steps, args.warmup, CPU_SAMPLE_BATCH)
    total = sum(times)
    value = CPU_SAMPLE_BATCH * len(times) / total
    sample = (f"{len(times)} steps of batch {CPU_SAMPLE_BATCH} (bounded sample of the batch-{args.batch} workload), "
              f"torch-CPU fp32 restatement of the TF/Larq path, {cores} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": len(times), "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload_name(args.batch), "sample_batch": CPU_SAMPLE_BATCH},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50",
                 "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, smax, power, reasons = [], [], [], set()
        for ln in self.f.read().splitlines():
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); smax.append(float(parts[1])); power.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(self.NAMES, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        # under load = samples drawing real power
        loaded = [s for s, p in zip(sm, power) if p > 300] or sm
        return {"sm_mhz": statistics.median(loaded) if loaded else None,
                "sm_max_mhz": max(smax) if smax else None, "power_w_max": max(power) if power else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# algorithmic work per launch
# ------------------------------------------------------------------------------------------------
def algorithmic_work(eng, name, tag):
    """('flop'|'byte', amount) for one launch of entry point `name` issued for layer `tag`."""
    g = global_work(eng, name)
    if g:
        return g
    i = int(tag[3:]) if tag[3:].isdigit() else None
    if i is None:
        return None
    L = eng.L[i]
    l = L.spec
    B = eng.B
    if l.kind == "conv":
        oh, ow = L.out_hw
        M, N, K = B * oh * ow, l.cout, l.ksize * l.ksize * l.cin
    else:
        M, N, K = B, l.cout, l.cin
    if name == "zkb_bconv2d_dgrad_first_bwd":
        # the dgrad GEMM plus the first layer's recomputed conv / BatchNorm-backward sums in its
        # epilogue (about 60 flop per output element: 9-tap conv, 2 + 18 accumulating FMAs)
        return ("flop", 2.0 * M * N * K + 60.0 * B * L.in_hw[0] * L.in_hw[1] * l.cin)
    if name in ("zkb_bconv2d_fwd", "zkb_bdense_fwd", "zkb_bconv2d_fwd_i8", "zkb_bdense_fwd_i8", "zkb_bconv2d_dgrad", "zkb_bdense_dgrad",
                "zkb_bconv2d_wgrad", "zkb_bdense_wgrad", "zkb_bconv2d_fwd_i8_stats", "zkb_bdense_fwd_i8_stats",
                "zkb_bconv2d_dgrad_bn"):
        return ("flop", 2.0 * M * N * K)  # split terms are implementation overhead, not counted
    rows, C = L.bn_rows, L.cp
    xb = 4 if i == 0 else 2
    table = {
        "zkb_bn_stats": rows * C * xb,
        "zkb_bn_apply_sign": rows * C * (xb + 2),
        "zkb_maxpool2_i16": (M * N * 2) + rows * C * 3,
        "zkb_bn_bwd_reduce": rows * C * (4 + xb),
        "zkb_bn_bwd_apply": rows * C * (4 + xb) + M * N * 4,
        "zkb_conv_first_bn_stats": B * 28 * 28 * 4,
        "zkb_conv_first_gram": B * 28 * 28 * 4,
        "zkb_conv_first_bn_sign": M * N * 2 + B * 28 * 28 * 4,
        "zkb_conv_first_bwd_reduce": M * N * 4 + B * 28 * 28 * 4,
        "zkb_conv_first_bwd_wgrad": M * N * 4 + B * 28 * 28 * 4,
    }
    if name in table:
        return ("byte", float(table[name]))
    return None


def global_work(eng, name):
    """Byte models of the launches that are not tied to one layer."""
    B = eng.B
    H, W, C = eng.input_shape
    if name == "zkb_preprocess_u8":
        return ("byte", float(B * H * W * C * 5 + B * 8))
    if name == "zkb_adam_clip":
        return ("byte", 28.0 * (eng.n_kernel + eng.n_beta) / 2)  # two launches: kernels, betas
    return None


_REAL_STDOUT = None


def _claim_stdout():
    """Libraries (NCCL prints its version banner) write to fd 1; the contract is ONE JSON line on
    stdout.  Route fd 1 to stderr for the whole run and keep the real stdout for the result line."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(line):
    _REAL_STDOUT.write(json.dumps(line) + "\n")
    _REAL_STDOUT.flush()


def main():
    args = parse()
    _claim_stdout()
    if args.impl == "reference":
        run_reference(args)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch multi-GPU runs with torch.distributed.run (one rank per GPU)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    B = args.batch
    n_pool = 8 * B
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    images = torch.randint(0, 256, (n_pool,) + INPUT_SHAPE, generator=g, dtype=torch.uint8)
    labels = torch.randint(0, NUM_CLASSES, (n_pool,), generator=g, dtype=torch.int64)
    d_images, d_labels = images.to(dev), labels.to(dev)
    index = [torch.arange(k * B, (k + 1) * B, dtype=torch.int32, device=dev) for k in range(8)]

    eng = BinaryNetEngine(binarynet_layers(INPUT_SHAPE, 128, 1024, NUM_CLASSES), INPUT_SHAPE, B,
                          NUM_CLASSES, learning_rate=5e-3, sync_bn=args.sync_bn)
    eng.init_glorot(0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    step_id = [0]

    graph_on = [False]

    def resident_step():
        k = step_id[0] % 8
        step_id[0] += 1
        if graph_on[0]:
            eng.train_step_graph(index[k])
        else:
            eng.train_step(d_images, d_labels, index[k], pad_size=PAD_SIZE, seed=1234 + rank)

    # ---- warm-up, then a one-step pass with every launch timed to find the dominant kernel
    for _ in range(max(args.warmup, 3)):
        resident_step()
    barrier()
    eng.watch, eng.watched = True, []
    resident_step()
    torch.cuda.synchronize()
    per = {}
    for name, tag, e0, e1 in eng.watched:
        per.setdefault((name, tag), []).append(e0.elapsed_time(e1))
    eng.watch, eng.watched = None, []
    dominant = max(per, key=lambda k: sum(per[k]))
    # per entry point: time, share of the (serial) step and achieved rate against its roofline
    peaks_hbm, peaks_tc = 6447.5, 1401.9
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            _p = json.load(f)
            peaks_hbm, peaks_tc = _p.get("hbm_gbs", peaks_hbm), _p.get("bf16_tflops_sustained", peaks_tc)
    except OSError:
        pass
    serial_ms = sum(sum(v) for v in per.values())
    kernels = {}
    for (name, tag), v in per.items():
        k = kernels.setdefault(name, {"ms": 0.0, "flop": 0.0, "byte": 0.0, "launches": 0})
        k["ms"] += sum(v)
        k["launches"] += len(v)
        w = algorithmic_work(eng, name, tag)
        if w:
            k[w[0]] += w[1] * len(v)
    kernel_table = {}
    for name, k in sorted(kernels.items(), key=lambda kv: -kv[1]["ms"]):
        if k["ms"] / serial_ms < 0.004:
            continue
        row = {"ms": round(k["ms"], 4), "share": round(k["ms"] / serial_ms, 4), "launches": k["launches"]}
        if k["flop"]:
            # int8 pipe: no measured peak in MEASURED_PEAKS.json; 2x the measured bf16 figure
            # (the nominal int8:bf16 ratio) is used and said so
            is_i8 = "_i8" in name
            pk = peaks_tc * (2.0 if is_i8 else 1.0)
            row.update({"achieved": round(k["flop"] / (k["ms"] * 1e-3) / 1e12, 1), "unit": "TFLOP/s", "bound": "tensor",
                        "peak": round(pk, 1), "frac": round(k["flop"] / (k["ms"] * 1e-3) / 1e12 / pk, 3)})
            if is_i8:
                row["peak_note"] = "2 x measured bf16 sustained"
        elif k["byte"]:
            row.update({"achieved": round(k["byte"] / (k["ms"] * 1e-3) / 1e9, 0), "unit": "GB/s", "bound": "hbm",
                        "frac": round(k["byte"] / (k["ms"] * 1e-3) / 1e9 / peaks_hbm, 3)})
        kernel_table[name] = row
    if args.breakdown and rank == 0:
        tot = sum(sum(v) for v in per.values())
        byname = {}
        for (name, tag), v in per.items():
            byname[name] = byname.get(name, 0.0) + sum(v)
        print(f"per-launch CUDA-event pass: {tot:.3f} ms in {len(eng.watched)} launches", file=sys.stderr)
        for (name, tag), v in sorted(per.items(), key=lambda kv: -sum(kv[1])):
            w = algorithmic_work(eng, name, tag)
            rate = ""
            if w:
                amt = w[1] / (sum(v) / len(v) * 1e-3)
                rate = f"{amt / 1e12:8.1f} TFLOP/s" if w[0] == "flop" else f"{amt / 1e9:8.0f} GB/s"
            print(f"  {name:24s} {tag:6s} {sum(v):8.3f} ms {100 * sum(v) / tot:5.1f}% {rate}", file=sys.stderr)
        for name, v in sorted(byname.items(), key=lambda kv: -kv[1]):
            print(f"  TOTAL {name:24s} {v:8.3f} ms {100 * v / tot:5.1f}%", file=sys.stderr)

    # ---- timed region (device-resident inputs)
    if args.graph:
        eng.capture_train_step(d_images, d_labels, PAD_SIZE, 1234 + rank)
        graph_on[0] = True
        for _ in range(3):
            resident_step()
    eng.watch = None if args.graph else {dominant[0]}
    clocks = ClockSampler(local_rank) if rank == 0 else None
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        resident_step()
    ev1.record()
    barrier()
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    clock_info = clocks.stop() if clocks else None
    launches = eng.launches_per_step * args.steps
    dom_ms = [e0.elapsed_time(e1) for name, tag, e0, e1 in eng.watched if (name, tag) == dominant]
    if args.graph:   # launches inside a replayed graph cannot carry events: use the serial pass
        dom_ms = per[dominant]
    graph_on[0] = False
    eng.watch, eng.watched = None, []
    value = world * B * args.steps / (ms_total * 1e-3)

    # ---- end-to-end: pinned host batches in, loss out, every step
    host_imgs = [images[k * B:(k + 1) * B].clone().pin_memory() for k in range(4)]
    host_labs = [labels[k * B:(k + 1) * B].clone().pin_memory() for k in range(4)]
    stage_img = torch.empty((B,) + INPUT_SHAPE, dtype=torch.uint8, device=dev)
    stage_lab = torch.empty((B,), dtype=torch.int64, device=dev)

    def host_step(k):
        stage_img.copy_(host_imgs[k % 4], non_blocking=True)
        stage_lab.copy_(host_labs[k % 4], non_blocking=True)
        eng.train_step(stage_img, stage_lab, None, pad_size=PAD_SIZE, seed=1234 + rank)
        return eng.last_loss()  # device->host read (synchronises)

    for k in range(3):
        host_step(k)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    loss = 0.0
    for k in range(args.steps):
        loss = host_step(k)
    e1.record()
    barrier()
    e2e_ms = max_over_ranks(e0.elapsed_time(e1))
    e2e_value = world * B * args.steps / (e2e_ms * 1e-3)
    h2d = B * int(np.prod(INPUT_SHAPE)) + B * 8
    d2h = 8

    # ---- gradient all-reduce in isolation (bus bandwidth = 2(n-1)/n * bytes / t)
    allreduce = None
    if world > 1:
        buf = eng.grads.clone()
        for _ in range(3):
            dist.all_reduce(buf)
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for _ in range(20):
            dist.all_reduce(buf)
        a1.record()
        barrier()
        ar_ms = max_over_ranks(a0.elapsed_time(a1)) / 20
        nbytes = buf.numel() * 4
        allreduce = {"bytes": nbytes, "ms": ar_ms, "bus_gbs": 2.0 * (world - 1) / world * nbytes / (ar_ms * 1e-3) / 1e9,
                     "backend": "nccl " + ".".join(str(v) for v in torch.cuda.nccl.version())}

    # ---- roofline of the dominant kernel
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    work = algorithmic_work(eng, *dominant)
    roofline = None
    if work and dom_ms:
        avg_s = statistics.mean(dom_ms) * 1e-3
        if work[0] == "flop":
            peak = peaks.get("bf16_tflops_sustained", 1400.0)
            ach = work[1] / avg_s / 1e12
            roofline = {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                        "traffic": None}
        else:
            peak = peaks.get("hbm_gbs", 6650.0)
            ach = work[1] / avg_s / 1e9
            roofline = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                        "traffic": None}
        try:
            with open(os.path.join(ROOT, "profiles", "r1_traffic.json")) as f:
                roofline["traffic"] = json.load(f).get(f"{dominant[0]}[{dominant[1]}]")
        except OSError:
            pass
        # the same launch timed alone (serial per-launch pass before the timed region): in the timed
        # region the step runs on two streams and this kernel shares the SMs with a wgrad GEMM
        ser_s = statistics.mean(per[dominant]) * 1e-3
        roofline["serial"] = {"avg_launch_ms": ser_s * 1e3, "achieved": work[1] / ser_s / (1e12 if work[0] == "flop" else 1e9),
                              "frac": work[1] / ser_s / (1e12 if work[0] == "flop" else 1e9) / roofline["peak"]}
        if work[0] == "flop" and "grad" in dominant[0]:
            roofline["note"] = ("algorithmic 2*M*N*K; the kernel issues twice that in bf16 MMAs (hi + lo split of dy, "
                                "needed for the 1e-4 parity budget), i.e. issued/peak = 2 x frac")
        roofline.update({"kernel": f"{dominant[0]}[{dominant[1]}]", "avg_launch_ms": statistics.mean(dom_ms),
                         "launches_timed": len(dom_ms),
                         "peak_source": "MEASURED_PEAKS.json (sustained)" if peaks else "fallback",
                         "share_of_step": statistics.mean(dom_ms) * len(dom_ms) / ms_total})

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(B), "global_batch": B * world, "parallelism": f"dp{world}",
                       "batchnorm": "sync" if eng.sync_bn else "per-replica",
                       "launch": "cuda-graph replay" if args.graph else "per-kernel launches, 2 streams",
                       "l2": "working set (GBs of activations per step) far larger than the 126 MB L2; no flush"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms / args.steps, "last_loss": loss},
            "gpu_launches": launches, "clocks": clock_info, "roofline": roofline,
            "kernels": kernel_table, "serial_ms_per_step": round(serial_ms, 3),
        }
        if allreduce:
            line["allreduce"] = allreduce
        if world == 1 and not args.no_cpu_baseline:
            times, cores = cpu_train_steps(40, 2, CPU_SAMPLE_BATCH, budget_s=20.0)
            v = CPU_SAMPLE_BATCH * len(times) / sum(times)
            line["cpu_baseline"] = {
                "value": v, "unit": UNIT, "cores": cores, "kind": "port",
                "sample": f"{len(times)} train steps of batch {CPU_SAMPLE_BATCH} (BASELINE.json configs[0]) of the "
                          f"torch-CPU fp32 restatement, {cores} threads"}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
