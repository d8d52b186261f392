"""CPU oracle for Bi-Real-Net style binary residual units -- TEST INFRASTRUCTURE ONLY
(imported by tests/ only; the product path never imports it).

PARITY UNPINNED.  The reference tree holds neither these operators nor a test for them:
BASELINE.json configs[3] / [4] name larq-zoo models (BinaryAlexNet, Bi-Real-Net-18), which a
zookeeper experiment reaches through the same hook as the example's BinaryNet -- a ``@factory`` model
built from larq layers with quantizer names in a ``kwhparams`` dict
(/root/reference/examples/larq_experiment.py:41-57, 59-103).  larq and larq-zoo are third-party
packages that are not vendored, not installed and not installable here (no network); the reference
does not even pin a larq version (.github/workflows/unittest.yml:37).  What follows restates their
PUBLISHED definitions (marked [EXT] there):

  larq.quantizers.approx_sign (Liu et al., Bi-Real Net):
      forward  sign(x) with sign(0) = +1
      backward dy * (2 - 2|x|) * [|x| <= 1]
  larq.quantizers.magnitude_aware_sign:
      scale = stop_gradient(mean(|w|) over every axis but the last)   (per output channel)
      q(w)  = scale * ste_sign(w, clip_value=1)                        (gradient dy * scale * [|w| <= 1])
  larq_zoo BiRealNet residual unit:
      h_next = BatchNormalization()(QuantConv2D(3x3, same, use_bias=False, input_quantizer=approx_sign,
               kernel_quantizer=magnitude_aware_sign, kernel_constraint=weight_clip)(h)) + h
      (Keras BatchNormalization defaults here: scale and centre, epsilon 1e-3; momentum does not
      enter a train step's gradients)
  head: GlobalAveragePooling2D -> Dense(num_classes) (float, with bias) -> softmax, sparse CE.
  stem pool: MaxPool2D(pool_size=3, strides=2, padding="same"); ties go to the first maximum in
      row-major window order (the same rule as the 2x2 pool).
"""
from __future__ import annotations

from typing import Dict, List

import numpy as np
import torch

BN_EPS = 1e-3


class _ApproxSign(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(x)
        return torch.where(x >= 0, torch.ones_like(x), -torch.ones_like(x))

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        a = x.abs()
        return dy * (2.0 - 2.0 * a) * (a <= 1).to(dy.dtype)


class _SteSign(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(x)
        return torch.where(x >= 0, torch.ones_like(x), -torch.ones_like(x))

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        return dy * (x.abs() <= 1).to(dy.dtype)


def approx_sign(x: torch.Tensor) -> torch.Tensor:
    return _ApproxSign.apply(x)


def magnitude_aware_sign(w_hwio: torch.Tensor) -> torch.Tensor:
    scale = w_hwio.detach().abs().mean(dim=(0, 1, 2))
    return scale * _SteSign.apply(w_hwio)


def init_params(units: int, channels: int, num_classes: int, seed: int = 0) -> Dict[str, np.ndarray]:
    """Glorot-uniform kernels, gamma 1 + noise, beta noise (so that no gradient is trivially zero)."""
    rng = np.random.default_rng(seed)
    p = {}
    for u in range(units):
        lim = np.sqrt(6.0 / (9 * channels + 9 * channels))
        p[f"kernel{u}"] = rng.uniform(-lim, lim, size=(3, 3, channels, channels)).astype(np.float32)
        p[f"gamma{u}"] = (1.0 + 0.1 * rng.standard_normal(channels)).astype(np.float32)
        p[f"beta{u}"] = (0.1 * rng.standard_normal(channels)).astype(np.float32)
    lim = np.sqrt(6.0 / (channels + num_classes))
    p["dense_kernel"] = rng.uniform(-lim, lim, size=(channels, num_classes)).astype(np.float32)
    p["dense_bias"] = (0.01 * rng.standard_normal(num_classes)).astype(np.float32)
    return p


def forward_backward(params: Dict[str, np.ndarray], units: int, h0: np.ndarray, labels: np.ndarray,
                     dtype=torch.float64, keep: bool = False):
    """One train-mode forward + backward of `units` residual units and the head.  h0: [B,H,W,C]
    float activations entering the first unit.  Returns dict(loss, logits, grads, acts)."""
    p = {k: torch.tensor(v, dtype=dtype, requires_grad=True) for k, v in params.items()}
    h = torch.tensor(h0, dtype=dtype, requires_grad=True)
    acts: Dict[str, np.ndarray] = {}
    x = h
    for u in range(units):
        s = approx_sign(x)
        wq = magnitude_aware_sign(p[f"kernel{u}"])
        z = torch.nn.functional.conv2d(s.permute(0, 3, 1, 2), wq.permute(3, 2, 0, 1), padding=1).permute(0, 2, 3, 1)
        mean = z.mean(dim=(0, 1, 2))
        var = z.var(dim=(0, 1, 2), unbiased=False)
        bn = (z - mean) * torch.rsqrt(var + BN_EPS) * p[f"gamma{u}"] + p[f"beta{u}"]
        if keep:
            with torch.no_grad():
                alpha = params[f"kernel{u}"].astype(np.float64)
                alpha = np.abs(alpha).mean(axis=(0, 1, 2))
                acts[f"y{u}"] = np.rint((z.detach().numpy() / alpha)).astype(np.int64)   # integer conv output
                acts[f"alpha{u}"] = alpha
                acts[f"mean_z{u}"], acts[f"var_z{u}"] = mean.detach().numpy(), var.detach().numpy()
        x = bn + x
        if keep:
            acts[f"h{u + 1}"] = x.detach().numpy()
    pooled = x.mean(dim=(1, 2))
    logits = pooled @ p["dense_kernel"] + p["dense_bias"]
    lab = torch.tensor(labels, dtype=torch.int64)
    loss = torch.nn.functional.cross_entropy(logits, lab)
    loss.backward()
    grads = {k: v.grad.detach().numpy() for k, v in p.items()}
    grads["h0"] = h.grad.detach().numpy()
    acc = float((logits.argmax(1) == lab).float().mean())
    return {"loss": float(loss.detach()), "logits": logits.detach().numpy(), "grads": grads, "acts": acts, "accuracy": acc}


def adam_step(params, grads, m, v, step: int, lr: float = 5e-3, b1=0.9, b2=0.999, eps=1e-7, dtype=np.float64):
    """tf.keras Adam (epsilon-hat form) on every parameter; latent kernels of the binary convs are
    then clipped to [-1, 1] (kernel_constraint="weight_clip")."""
    f = dtype
    lr_t = f(lr) * np.sqrt(f(1) - f(b2) ** f(step), dtype=f) / (f(1) - f(b1) ** f(step))
    out, mo, vo = {}, {}, {}
    for k, w in params.items():
        g = grads[k].astype(f)
        mm = m[k].astype(f) + (g - m[k].astype(f)) * (f(1) - f(b1))
        vv = v[k].astype(f) + (g * g - v[k].astype(f)) * (f(1) - f(b2))
        nw = w.astype(f) - (mm * lr_t) / (np.sqrt(vv) + f(eps))
        if k.startswith("kernel"):
            nw = np.clip(nw, -1.0, 1.0)
        out[k], mo[k], vo[k] = nw.astype(np.float32), mm.astype(np.float32), vv.astype(np.float32)
    return out, mo, vo


def maxpool3s2_same(x: np.ndarray):
    """MaxPool2D(3, 2, 'same') on float NHWC; returns (out, slot) with slot = kh*3+kw of the first
    maximum in row-major window order."""
    b, h, w, c = x.shape
    oh, ow = (h + 1) // 2, (w + 1) // 2
    ph, pw = max((oh - 1) * 2 + 3 - h, 0) // 2, max((ow - 1) * 2 + 3 - w, 0) // 2
    out = np.full((b, oh, ow, c), -np.inf, dtype=x.dtype)
    slot = np.zeros((b, oh, ow, c), dtype=np.uint8)
    for kh in range(3):
        for kw in range(3):
            for i in range(oh):
                ih = 2 * i + kh - ph
                if ih < 0 or ih >= h:
                    continue
                for j in range(ow):
                    iw = 2 * j + kw - pw
                    if iw < 0 or iw >= w:
                        continue
                    v = x[:, ih, iw, :]
                    better = v > out[:, i, j, :]
                    out[:, i, j, :] = np.where(better, v, out[:, i, j, :])
                    slot[:, i, j, :] = np.where(better, kh * 3 + kw, slot[:, i, j, :])
    return out, slot


def maxpool3s2_same_backward(dout: np.ndarray, slot: np.ndarray, in_hw) -> np.ndarray:
    b, oh, ow, c = dout.shape
    h, w = in_hw
    ph, pw = max((oh - 1) * 2 + 3 - h, 0) // 2, max((ow - 1) * 2 + 3 - w, 0) // 2
    dx = np.zeros((b, h, w, c), dtype=dout.dtype)
    for i in range(oh):
        for j in range(ow):
            kh, kw = slot[:, i, j, :] // 3, slot[:, i, j, :] % 3
            ih, iw = 2 * i + kh - ph, 2 * j + kw - pw
            bi, ci = np.meshgrid(np.arange(b), np.arange(c), indexing="ij")
            np.add.at(dx, (bi, ih, iw, ci), dout[:, i, j, :])
    return dx
