"""CPU oracle for a BinaryAlexNet train step -- TEST INFRASTRUCTURE ONLY (imported by tests/ and, as the
timed CPU baseline only, by scripts/bench_zoo.py --cpu-baseline; the product path never imports it).

PARITY UNPINNED.  BASELINE.json configs[3] names larq-zoo's BinaryAlexNet, which a zookeeper experiment
reaches through the same hook as the example's BinaryNet: a ``@factory`` model built from larq layers
with the quantizer names in a ``kwhparams`` dict (/root/reference/examples/larq_experiment.py:41-57,
59-103; compiled and fitted at :117-122, :141-153).  larq-zoo is a third-party package that is not
vendored, not installed and not installable here; the reference pins no version of it.  What follows
restates its PUBLISHED definition (marked [EXT] there) on top of the operator
semantics oracle/binarynet_oracle.py already restates (ste_sign, BatchNormalization(scale=False),
Adam + weight_clip):

  conv_block(features, kernel, strides=1, pool=False, first=False):
      QuantConv2D(features, kernel, strides, padding="same", use_bias=False,
                  input_quantizer=None if first else "ste_sign", kernel_quantizer="ste_sign",
                  kernel_constraint="weight_clip")
      -> MaxPool2D(pool_size=3, strides=2) if pool        ("valid": out = (in - 3) // 2 + 1)
      -> BatchNormalization(scale=False, momentum=0.9)
  dense_block(units): QuantDense(units, ste_sign / ste_sign / weight_clip, use_bias=False)
      -> BatchNormalization(scale=False, momentum=0.9)
  model: conv_block(64, 11, strides=4, pool, first) -> conv_block(192, 5, pool) -> conv_block(384, 3)
      -> conv_block(384, 3) -> conv_block(256, 3, pool) -> Flatten -> dense_block(4096) x 2
      -> dense_block(num_classes) -> softmax; sparse categorical cross-entropy; Adam.
  MaxPool2D ties: the first maximum in row-major window order takes the gradient (SURVEY.md
  Appendix A.6) -- the pooled maps are integers, so ties are common.
"""
from __future__ import annotations

from typing import Dict, Sequence, Tuple

import numpy as np
import torch

from .binarynet_oracle import BN_EPS, adam_clip_step, ste_sign  # noqa: F401  (adam_clip_step re-exported)

FEATURES = (64, 192, 384, 384, 256)
KSIZES = (11, 5, 3, 3, 3)
STRIDES = (4, 1, 1, 1, 1)
POOLS = (True, True, False, False, True)


def shapes(image_hw: Tuple[int, int], features: Sequence[int] = FEATURES):
    """[(in_h, in_w, cin, conv_h, conv_w, cout, out_h, out_w)] per conv block, and the flat width."""
    h, w, c = image_hw[0], image_hw[1], 3
    out = []
    for f, s, pool in zip(features, STRIDES, POOLS):
        ch, cw = -(-h // s), -(-w // s)
        oh, ow = ((ch - 3) // 2 + 1, (cw - 3) // 2 + 1) if pool else (ch, cw)
        if oh < 1 or ow < 1:
            raise ValueError("image too small for the three 3x3/2 pools")
        out.append((h, w, c, ch, cw, f, oh, ow))
        h, w, c = oh, ow, f
    return out, h * w * c


def init_params(num_classes: int, image_hw=(224, 224), features: Sequence[int] = FEATURES, dense_units: int = 4096,
                seed: int = 0) -> Dict[str, np.ndarray]:
    """Glorot-uniform latent kernels (``kernel0..7``) and small random betas (``beta0..7``)."""
    rng = np.random.default_rng(seed)
    p = {}
    sh, flat = shapes(image_hw, features)
    for i, ((_, _, cin, _, _, cout, _, _), k) in enumerate(zip(sh, KSIZES)):
        lim = np.sqrt(6.0 / (k * k * (cin + cout)))
        p[f"kernel{i}"] = rng.uniform(-lim, lim, size=(k, k, cin, cout)).astype(np.float32)
        p[f"beta{i}"] = (0.1 * rng.standard_normal(cout)).astype(np.float32)
    for j, (kin, kout) in enumerate(((flat, dense_units), (dense_units, dense_units), (dense_units, num_classes))):
        lim = np.sqrt(6.0 / (kin + kout))
        p[f"kernel{5 + j}"] = rng.uniform(-lim, lim, size=(kin, kout)).astype(np.float32)
        p[f"beta{5 + j}"] = (0.1 * rng.standard_normal(kout)).astype(np.float32)
    return p


class _MaxPool3S2First(torch.autograd.Function):
    """MaxPool2D(3, 2) 'valid' on NHWC; the running maximum is replaced on strict '>' only."""

    @staticmethod
    def forward(ctx, x):
        n, h, w, c = x.shape
        oh, ow = (h - 3) // 2 + 1, (w - 3) // 2 + 1
        best = torch.full((n, oh, ow, c), float("-inf"), dtype=x.dtype)
        arg = torch.zeros((n, oh, ow, c), dtype=torch.int64)
        for kh in range(3):
            for kw in range(3):
                v = x[:, kh:kh + 2 * oh - 1:2, kw:kw + 2 * ow - 1:2]
                better = v > best
                best = torch.where(better, v, best)
                arg = torch.where(better, torch.full_like(arg, kh * 3 + kw), arg)
        ctx.save_for_backward(arg)
        ctx.in_shape = x.shape
        return best

    @staticmethod
    def backward(ctx, dy):
        (arg,) = ctx.saved_tensors
        n, h, w, c = ctx.in_shape
        oh, ow = dy.shape[1:3]
        dx = torch.zeros(n, h, w, c, dtype=dy.dtype)
        for kh in range(3):
            for kw in range(3):
                dx[:, kh:kh + 2 * oh - 1:2, kw:kw + 2 * ow - 1:2] += dy * (arg == kh * 3 + kw).to(dy.dtype)
        return dx


def maxpool3s2_valid(x: torch.Tensor) -> torch.Tensor:
    return _MaxPool3S2First.apply(x)


def _same_pad(x_nchw, k, s):
    h, w = x_nchw.shape[2:]
    th = max((-(-h // s) - 1) * s + k - h, 0)
    tw = max((-(-w // s) - 1) * s + k - w, 0)
    return torch.nn.functional.pad(x_nchw, (tw // 2, tw - tw // 2, th // 2, th - th // 2))


def _bn(z, beta, axes):
    mean = z.mean(dim=axes)
    var = z.var(dim=axes, unbiased=False)
    return (z - mean) * torch.rsqrt(var + BN_EPS) + beta


def forward_backward(params: Dict[str, np.ndarray], images: np.ndarray, labels: np.ndarray, dtype=torch.float64,
                     keep: bool = False):
    """One train-mode forward + backward.  images: float [B,H,W,3] (already scaled to [-1, 1]).
    Returns dict(loss, logits, grads, acts, accuracy); acts (keep=True): ``y{i}`` the conv / dense
    outputs before pooling (integers from block 1 on), ``p{i}`` the pooled maps."""
    p = {k: torch.tensor(v, dtype=dtype, requires_grad=True) for k, v in params.items()}
    x = torch.tensor(images, dtype=dtype)
    acts = {}
    for i, (k, s, pool) in enumerate(zip(KSIZES, STRIDES, POOLS)):
        a = x if i == 0 else ste_sign(x)
        wq = ste_sign(p[f"kernel{i}"])
        y = torch.nn.functional.conv2d(_same_pad(a.permute(0, 3, 1, 2), k, s), wq.permute(3, 2, 0, 1), stride=s).permute(0, 2, 3, 1)
        if keep:
            acts[f"y{i}"] = y.detach().numpy()
        if pool:
            y = maxpool3s2_valid(y)
            if keep:
                acts[f"p{i}"] = y.detach().numpy()
        x = _bn(y, p[f"beta{i}"], (0, 1, 2))
    x = x.reshape(x.shape[0], -1)                      # Flatten: (h, w, c) order of NHWC
    for j in range(5, 8):
        y = ste_sign(x) @ ste_sign(p[f"kernel{j}"])
        if keep:
            acts[f"y{j}"] = y.detach().numpy()
        x = _bn(y, p[f"beta{j}"], (0,))
    logits = x
    lab = torch.tensor(labels, dtype=torch.int64)
    loss = torch.nn.functional.cross_entropy(logits, lab)
    loss.backward()
    grads = {k: v.grad.detach().numpy() for k, v in p.items()}
    acc = float((logits.argmax(1) == lab).float().mean())
    return {"loss": float(loss.detach()), "logits": logits.detach().numpy(), "grads": grads, "acts": acts, "accuracy": acc}
