"""GPU probe for the tcgen05 GEMM engines: one case per process (a trap poisons the context).

usage: python scripts/probe_gemm.py <case>   |   python scripts/probe_gemm.py --list
"""
import sys

import torch

sys.path.insert(0, ".")
from zookeeper_b200 import native as zn  # noqa: E402

dev = "cuda"


def pm1(*shape, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randint(0, 2, shape, generator=g, dtype=torch.int8) * 2 - 1
    return x.to(dev)


def split(x):
    hi = x.to(torch.bfloat16)
    lo = (x - hi.float()).to(torch.bfloat16)
    return hi.contiguous(), lo.contiguous()


def report(name, got, ref, exact=False):
    got = got.double()
    ref = ref.double()
    err = (got - ref).abs()
    rel = err.norm() / max(ref.norm().item(), 1e-30)
    ok = (err.max().item() == 0) if exact else (rel.item() < 2e-5)
    print(f"[{name}] max_abs_err={err.max().item():.4g} rel_fro={rel.item():.3g} ref_norm={ref.norm().item():.4g} ok={ok}")
    if not ok:
        bad = (err > (0 if exact else 1e-3 * ref.abs().max())).nonzero()
        print("  n_bad", bad.shape[0], "of", err.numel(), "first bad idx", bad[:8].tolist())
        flat = err.reshape(err.shape[0], -1) if err.dim() > 1 else err.reshape(1, -1)
        rows_bad = (flat.max(dim=1).values > 0).nonzero().flatten()
        print("  bad rows (first 16):", rows_bad[:16].tolist(), "count", rows_bad.numel())
        cols_bad = (flat.max(dim=0).values > 0).nonzero().flatten()
        print("  bad cols (first 16):", cols_bad[:16].tolist(), "count", cols_bad.numel())
        print("  got[0,:8]", got.reshape(got.shape[0], -1)[0, :8].tolist())
        print("  ref[0,:8]", ref.reshape(ref.shape[0], -1)[0, :8].tolist())
    return ok


def dense_fwd(M, K, N):
    a = pm1(M, K, seed=1)
    w = pm1(N, K, seed=2)
    y = torch.full((M, N), -7, dtype=torch.int16, device=dev)
    ab, wb = a.to(torch.bfloat16).contiguous(), w.to(torch.bfloat16).contiguous()  # keep alive
    zn.call("zkb_bdense_fwd", zn.ptr(ab), zn.ptr(wb), zn.ptr(y), N, M, K, N, 0)
    torch.cuda.synchronize()
    ref = a.float() @ w.float().t()
    return report(f"dense_fwd M={M} K={K} N={N}", y, ref, exact=True)


def conv_ref(x_nhwc, w_hwio):
    # 'same' zero-padded correlation, NHWC / HWIO like Keras
    x = x_nhwc.permute(0, 3, 1, 2).double()
    w = w_hwio.permute(3, 2, 0, 1).double()
    y = torch.nn.functional.conv2d(x, w, padding=w_hwio.shape[0] // 2)
    return y.permute(0, 2, 3, 1).contiguous()


def prep_conv(w_hwio):
    ks, _, cin, cout = w_hwio.shape
    wf = torch.empty(cout, ks * ks * cin, dtype=torch.bfloat16, device=dev)
    wd = torch.empty(cin, ks * ks * cout, dtype=torch.bfloat16, device=dev)
    zn.call("zkb_prep_conv_weights", zn.ptr(w_hwio.float().contiguous()), ks, cin, cout, zn.ptr(wf), zn.ptr(wd), 0, 0)
    return wf, wd


def conv_fwd(B, H, W, Cin, Cout):
    x = pm1(B, H, W, Cin, seed=3)
    w = pm1(3, 3, Cin, Cout, seed=4)
    wf, _ = prep_conv(w)
    y = torch.full((B, H, W, Cout), -7, dtype=torch.int16, device=dev)
    zn.call("zkb_bconv2d_fwd", zn.ptr(x.to(torch.bfloat16).contiguous()), zn.ptr(wf), zn.ptr(y), B, H, W, Cin, Cout, 3, 0)
    torch.cuda.synchronize()
    return report(f"conv_fwd B={B} {H}x{W} {Cin}->{Cout}", y, conv_ref(x, w), exact=True)


def conv_dgrad(B, H, W, Cin, Cout):
    g = torch.Generator(device="cpu").manual_seed(5)
    dy = (torch.randn(B, H, W, Cout, generator=g) * torch.rand(B, H, W, Cout, generator=g)).to(dev)
    w = pm1(3, 3, Cin, Cout, seed=6)
    _, wd = prep_conv(w)
    hi, lo = split(dy)
    dx = torch.full((B, H, W, Cin), 7.0, dtype=torch.float32, device=dev)
    zn.call("zkb_bconv2d_dgrad", zn.ptr(hi), zn.ptr(lo), zn.ptr(wd), zn.ptr(dx), B, H, W, Cin, Cout, 3, 0)
    torch.cuda.synchronize()
    x = torch.zeros(B, Cin, H, W, dtype=torch.double, device=dev, requires_grad=True)
    y = torch.nn.functional.conv2d(x, w.permute(3, 2, 0, 1).double(), padding=1)
    y.backward(dy.permute(0, 3, 1, 2).double())
    ref = x.grad.permute(0, 2, 3, 1)
    return report(f"conv_dgrad B={B} {H}x{W} {Cin}<-{Cout}", dx, ref)


def conv_wgrad(B, H, W, Cin, Cout):
    g = torch.Generator(device="cpu").manual_seed(7)
    dy = torch.randn(B, H, W, Cout, generator=g).to(dev)
    x = pm1(B, H, W, Cin, seed=8)
    hi, lo = split(dy)
    dw = torch.zeros(3, 3, Cin, Cout, dtype=torch.float32, device=dev)
    zn.call("zkb_bconv2d_wgrad", zn.ptr(x.to(torch.bfloat16).contiguous()), zn.ptr(hi), zn.ptr(lo), zn.ptr(dw),
            B, H, W, Cin, Cout, 3, 0)
    torch.cuda.synchronize()
    w = torch.zeros(Cout, Cin, 3, 3, dtype=torch.double, device=dev, requires_grad=True)
    y = torch.nn.functional.conv2d(x.permute(0, 3, 1, 2).double(), w, padding=1)
    y.backward(dy.permute(0, 3, 1, 2).double())
    ref = w.grad.permute(2, 3, 1, 0)
    return report(f"conv_wgrad B={B} {H}x{W} {Cin}x{Cout}", dw.reshape(9 * Cin, Cout), ref.reshape(9 * Cin, Cout))


def dense_dgrad(M, K, N):
    g = torch.Generator(device="cpu").manual_seed(9)
    dy = torch.randn(M, N, generator=g).to(dev)
    w = pm1(K, N, seed=10)
    hi, lo = split(dy)
    dx = torch.full((M, K), 7.0, dtype=torch.float32, device=dev)
    zn.call("zkb_bdense_dgrad", zn.ptr(hi), zn.ptr(lo), zn.ptr(w.to(torch.bfloat16).contiguous()), zn.ptr(dx), M, K, N, 0)
    torch.cuda.synchronize()
    return report(f"dense_dgrad M={M} K={K} N={N}", dx, dy.double() @ w.double().t())


def dense_wgrad(M, K, N):
    g = torch.Generator(device="cpu").manual_seed(11)
    dy = torch.randn(M, N, generator=g).to(dev)
    x = pm1(M, K, seed=12)
    hi, lo = split(dy)
    dw = torch.zeros(K, N, dtype=torch.float32, device=dev)
    zn.call("zkb_bdense_wgrad", zn.ptr(x.to(torch.bfloat16).contiguous()), zn.ptr(hi), zn.ptr(lo), zn.ptr(dw), N, M, K, N, N, 0)
    torch.cuda.synchronize()
    return report(f"dense_wgrad M={M} K={K} N={N}", dw, x.double().t() @ dy.double())


CASES = {
    "dense_fwd_small": lambda: dense_fwd(128, 64, 128),
    "dense_fwd_k": lambda: dense_fwd(128, 1024, 128),
    "dense_fwd_ragged": lambda: dense_fwd(300, 1024, 64),
    "dense_fwd_big": lambda: dense_fwd(4096, 4608, 1024),
    "conv_fwd_tiny": lambda: conv_fwd(1, 6, 6, 128, 128),
    "conv_fwd_26": lambda: conv_fwd(3, 26, 26, 128, 128),
    "conv_fwd_13": lambda: conv_fwd(5, 13, 13, 256, 256),
    "conv_fwd_6": lambda: conv_fwd(7, 6, 6, 512, 512),
    "conv_dgrad_13": lambda: conv_dgrad(5, 13, 13, 128, 256),
    "conv_dgrad_26": lambda: conv_dgrad(3, 26, 26, 128, 128),
    "conv_wgrad_13": lambda: conv_wgrad(5, 13, 13, 128, 256),
    "conv_wgrad_26": lambda: conv_wgrad(3, 26, 26, 128, 128),
    "conv_wgrad_6": lambda: conv_wgrad(64, 6, 6, 512, 512),
    "dense_dgrad": lambda: dense_dgrad(256, 4608, 1024),
    "dense_dgrad_64": lambda: dense_dgrad(256, 1024, 64),
    "dense_wgrad": lambda: dense_wgrad(512, 1024, 1024),
    "dense_wgrad_64": lambda: dense_wgrad(512, 1024, 64),
}

if __name__ == "__main__":
    if sys.argv[1] == "--list":
        print(" ".join(CASES))
    else:
        ok = CASES[sys.argv[1]]()
        sys.exit(0 if ok else 1)
