"""Stand-alone timing of the GEMM entry points at the BinaryNet layer shapes (batch 4096)."""
import os
import sys

import torch

sys.path.insert(0, ".")
from zookeeper_b200 import native as zn  # noqa: E402

B = int(os.environ.get("B", "4096"))
LAYERS = [(26, 26, 128, 128), (13, 13, 128, 256), (13, 13, 256, 256), (6, 6, 256, 512), (6, 6, 512, 512)]
which = sys.argv[1:] or ["fwd", "fwd8", "dgrad", "wgrad"]


ONCE = os.environ.get("GEMM_ONCE") == "1"      # one launch per kernel (profiling runs)


def timeit(fn, n=8):
    if ONCE:
        n = 1
    for _ in range(0 if ONCE else 2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for (H, W, Cin, Cout) in LAYERS:
    M, K = B * H * W, 9 * Cin
    s = (torch.randint(0, 2, (B, H, W, Cin), device="cuda", dtype=torch.int8) * 2 - 1)
    sb = s.to(torch.bfloat16)
    w = torch.randn(3, 3, Cin, Cout, device="cuda")
    wf = torch.empty(Cout, K, dtype=torch.bfloat16, device="cuda")
    wd = torch.empty(Cin, 9 * Cout, dtype=torch.bfloat16, device="cuda")
    wf8 = torch.empty(Cout, K, dtype=torch.int8, device="cuda")
    zn.call("zkb_prep_conv_weights", zn.ptr(w), 3, Cin, Cout, zn.ptr(wf), zn.ptr(wd), zn.ptr(wf8), 0)
    y = torch.empty(B, H, W, Cout, dtype=torch.int16, device="cuda")
    dy = torch.randn(B, H, W, Cout, device="cuda")
    hi = dy.to(torch.bfloat16)
    lo = (dy - hi.float()).to(torch.bfloat16)
    dx = torch.empty(B, H, W, Cin, dtype=torch.float32, device="cuda")
    dw = torch.zeros(3, 3, Cin, Cout, device="cuda")
    flop = 2.0 * M * Cout * K
    out = [f"{H}x{W} {Cin}->{Cout}"]
    if "fwd" in which:
        t = timeit(lambda: zn.call("zkb_bconv2d_fwd", zn.ptr(sb), zn.ptr(wf), zn.ptr(y), B, H, W, Cin, Cout, 3, 0))
        out.append(f"fwd {t:.3f} ms {flop / t / 1e9:.0f} TF/s")
    if "fwd8" in which:
        t = timeit(lambda: zn.call("zkb_bconv2d_fwd_i8", zn.ptr(s), zn.ptr(wf8), zn.ptr(y), B, H, W, Cin, Cout, 3, 0))
        out.append(f"fwd8 {t:.3f} ms {flop / t / 1e9:.0f} TF/s")
    if "dgrad" in which:
        t = timeit(lambda: zn.call("zkb_bconv2d_dgrad", zn.ptr(hi), zn.ptr(lo), zn.ptr(wd), zn.ptr(dx), B, H, W, Cin, Cout, 3, 0))
        out.append(f"dgrad {t:.3f} ms {flop / t / 1e9:.0f} TF/s")
    if "wgrad" in which:
        t = timeit(lambda: zn.call("zkb_bconv2d_wgrad", zn.ptr(sb), zn.ptr(hi), zn.ptr(lo), zn.ptr(dw), B, H, W, Cin, Cout, 3, 0))
        out.append(f"wgrad {t:.3f} ms {flop / t / 1e9:.0f} TF/s")
    print(" | ".join(out), flush=True)
