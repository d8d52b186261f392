"""Not a test: how many mantissa bits the backward GEMM operand dy needs (run by hand on CPU).

The oracle's f64 backward is repeated with dy (the gradient w.r.t. each conv/dense output, i.e. the
float operand of the dgrad / wgrad GEMMs) rounded the way a given tensor-core operand split would
round it; prints the worst per-tensor Frobenius error of the gradients against the unrounded run.
    python tests/precision_study.py [batch]
"""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from oracle import binarynet_oracle as O  # noqa: E402


def q_bf16x2(g):
    hi = g.to(torch.bfloat16).to(g.dtype)
    return hi + (g - hi).to(torch.bfloat16).to(g.dtype)


def q_bf16(g):
    return g.to(torch.bfloat16).to(g.dtype)


def q_fp16(g):
    s = 2.0 ** (13 - np.ceil(np.log2(float(g.abs().max()) + 1e-300)))
    return (g * s).to(torch.float16).to(g.dtype) / s


def _e4m3(x):
    return x.to(torch.float32).to(torch.float8_e4m3fn).to(torch.float64)


def q_fp16_e4m3(g):
    s = 2.0 ** (13 - np.ceil(np.log2(float(g.abs().max()) + 1e-300)))
    v = g * s
    hi = v.to(torch.float16).to(g.dtype)
    lo = _e4m3((v - hi) * 64.0) / 64.0
    return (hi + lo) / s


def q_bf16_e4m3(g):
    s = 2.0 ** (13 - np.ceil(np.log2(float(g.abs().max()) + 1e-300)))
    v = g * s
    hi = v.to(torch.bfloat16).to(g.dtype)
    lo = _e4m3((v - hi) * 8.0) / 8.0
    return (hi + lo) / s


class _RoundGrad(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, fn):
        ctx.fn = fn
        return x.view_as(x)

    @staticmethod
    def backward(ctx, dy):
        return ctx.fn(dy), None


def run(params, layers, x, labels, fn):
    p = {k: torch.tensor(v, dtype=torch.float64, requires_grad=True) for k, v in params.items()}
    h = torch.tensor(x, dtype=torch.float64)
    flat = False
    for i, l in enumerate(layers):
        if l.kind == "dense" and not flat:
            h, flat = h.reshape(h.shape[0], -1), True
        a_in = O.ste_sign(h) if l.quant_input else h
        wq = O.ste_sign(p[f"kernel{i}"])
        y = O._conv_nhwc(a_in, wq, l.padding) if l.kind == "conv" else a_in @ wq
        if fn is not None:
            y = _RoundGrad.apply(y, fn)
        if l.pool:
            y = O.maxpool2(y)
        red = tuple(range(y.dim() - 1))
        mean = y.mean(dim=red)
        var = ((y - mean) ** 2).mean(dim=red)
        h = (y - mean) * torch.rsqrt(var + O.BN_EPS) + p[f"beta{i}"]
    loss = torch.nn.functional.cross_entropy(h, torch.tensor(labels, dtype=torch.int64))
    loss.backward()
    return {k: v.grad.numpy().copy() for k, v in p.items()}


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    layers = O.binarynet_layers((28, 28, 1), 128, 1024, 10)
    params = O.glorot_uniform_params(layers, 0)
    rng = np.random.default_rng(1234)
    img = rng.integers(0, 256, (B, 28, 28, 1), dtype=np.uint8)
    lab = rng.integers(0, 10, (B,)).astype(np.int64)
    x = O.preprocess(img, (28, 28), True, 32, O.augmentation_params(1234, 0, B, 32, 28, 28))
    ref = run(params, layers, x, lab, None)
    for name, fn in [("bf16", q_bf16), ("fp16", q_fp16), ("bf16+bf16 (current)", q_bf16x2),
                     ("fp16+e4m3", q_fp16_e4m3), ("bf16+e4m3", q_bf16_e4m3)]:
        got = run(params, layers, x, lab, fn)
        errs = {k: np.linalg.norm(got[k] - ref[k]) / max(np.linalg.norm(ref[k]), 1e-300) for k in ref}
        worst = max(errs, key=errs.get)
        print(f"{name:22s} worst {errs[worst]:.3g} ({worst})   median {np.median(list(errs.values())):.3g}")


if __name__ == "__main__":
    main()
