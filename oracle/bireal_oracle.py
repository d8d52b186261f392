"""CPU oracle for Bi-Real-Net style binary residual units -- TEST INFRASTRUCTURE ONLY
(imported by tests/ and, as the timed CPU baseline only, by scripts/bench_zoo.py --cpu-baseline; the
product path never imports it).

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
  larq_zoo BiRealNet down-sampling unit (first unit of stages 2-4):
      main     = BatchNormalization()(QuantConv2D(3x3, strides=2, same, ... as above)(h))   (C -> 2C)
      shortcut = BatchNormalization()(Conv2D(2C, 1, use_bias=False)(AveragePooling2D(2, 2)(h)))   (float)
      h_next   = main + shortcut
      (TensorFlow 'same' at stride 2: out = ceil(in/2), total padding max((out-1)*2 + 3 - in, 0), the
      smaller half in front -- even maps pad 0 before and 1 after)
  head: GlobalAveragePooling2D -> Dense(num_classes) (float, with bias) -> softmax, sparse CE.
  stem pool: MaxPool2D(pool_size=3, strides=2, padding="same"); ties go to the first maximum in
      row-major window order (the same rule as the 2x2 pool).
  larq_zoo BiRealNet stem: Conv2D(64, 7, strides=2, padding="same", use_bias=False) on the image (float
      weights, no quantizer, no constraint) -> BatchNormalization() -> the stem pool; then four
      64-channel units and three stages that open with a down-sampling unit (``bireal18_plan``).
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


def _plan(units):
    return ["plain"] * units if isinstance(units, int) else list(units)


def bireal18_plan():
    return ["plain"] * 4 + (["down"] + ["plain"] * 3) * 3


def init_params(units, channels: int, num_classes: int, seed: int = 0, stem_ksize: int = 0) -> Dict[str, np.ndarray]:
    """Glorot-uniform kernels, gamma 1 + noise, beta noise (so that no gradient is trivially zero).
    ``units``: a count of plain units or a sequence of "plain" / "down".  ``stem_ksize`` > 0 adds the
    stem's parameters (``channels`` filters on a 3-channel image)."""
    rng = np.random.default_rng(seed)
    p = {}
    c = channels
    if stem_ksize:
        lim = np.sqrt(6.0 / (stem_ksize * stem_ksize * (3 + c)))
        p["stem_kernel"] = rng.uniform(-lim, lim, size=(stem_ksize, stem_ksize, 3, c)).astype(np.float32)
        p["stem_gamma"] = (1.0 + 0.1 * rng.standard_normal(c)).astype(np.float32)
        p["stem_beta"] = (0.1 * rng.standard_normal(c)).astype(np.float32)
    for u, kind in enumerate(_plan(units)):
        co = 2 * c if kind == "down" else c
        lim = np.sqrt(6.0 / (9 * c + 9 * co))
        p[f"kernel{u}"] = rng.uniform(-lim, lim, size=(3, 3, c, co)).astype(np.float32)
        p[f"gamma{u}"] = (1.0 + 0.1 * rng.standard_normal(co)).astype(np.float32)
        p[f"beta{u}"] = (0.1 * rng.standard_normal(co)).astype(np.float32)
        if kind == "down":
            lim = np.sqrt(6.0 / (c + co))
            p[f"sc_kernel{u}"] = rng.uniform(-lim, lim, size=(c, co)).astype(np.float32)
            p[f"sc_gamma{u}"] = (1.0 + 0.1 * rng.standard_normal(co)).astype(np.float32)
            p[f"sc_beta{u}"] = (0.1 * rng.standard_normal(co)).astype(np.float32)
        c = co
    lim = np.sqrt(6.0 / (c + num_classes))
    p["dense_kernel"] = rng.uniform(-lim, lim, size=(c, num_classes)).astype(np.float32)
    p["dense_bias"] = (0.01 * rng.standard_normal(num_classes)).astype(np.float32)
    return p


def _batchnorm(z, gamma, beta):
    mean = z.mean(dim=(0, 1, 2))
    var = z.var(dim=(0, 1, 2), unbiased=False)
    return (z - mean) * torch.rsqrt(var + BN_EPS) * gamma + beta, mean, var


def _same_pad(x_nchw, k, s, value=0.0):
    """TensorFlow 'same': out = ceil(in / s), total padding max((out-1) s + k - in, 0), smaller half first."""
    h, w = x_nchw.shape[2:]
    th = max((-(-h // s) - 1) * s + k - h, 0)
    tw = max((-(-w // s) - 1) * s + k - w, 0)
    return torch.nn.functional.pad(x_nchw, (tw // 2, tw - tw // 2, th // 2, th - th // 2), value=value)


def _same_pad_stride2(x_nchw, k=3):
    return _same_pad(x_nchw, k, 2)


def stem(p, images, stride: int = 2):
    """Conv2D(k, strides, 'same', no bias) -> BatchNormalization -> MaxPool2D(3, 2, 'same') on NHWC
    images; returns (h [B,H2,W2,C], conv output z, BatchNorm output a)."""
    k = p["stem_kernel"].shape[0]
    z = torch.nn.functional.conv2d(_same_pad(images.permute(0, 3, 1, 2), k, stride), p["stem_kernel"].permute(3, 2, 0, 1),
                                   stride=stride).permute(0, 2, 3, 1)
    a, _, _ = _batchnorm(z, p["stem_gamma"], p["stem_beta"])
    # -inf padding: a padded position never wins; torch's CPU pool keeps the first maximum in scan order
    h = torch.nn.functional.max_pool2d(_same_pad(a.permute(0, 3, 1, 2), 3, 2, value=float("-inf")), 3, 2).permute(0, 2, 3, 1)
    return h, z, a


def forward_backward(params: Dict[str, np.ndarray], units, h0: np.ndarray, labels: np.ndarray,
                     dtype=torch.float64, keep: bool = False):
    """One train-mode forward + backward of the residual units and the head.  h0: [B,H,W,C] float
    activations entering the first unit -- or, when ``params`` holds ``stem_kernel``, the float images
    [B,H,W,3] entering the stem (``stem_stride`` from the kernel size: 7 -> 2, 11 -> 4).
    Returns dict(loss, logits, grads, acts)."""
    p = {k: torch.tensor(v, dtype=dtype, requires_grad=True) for k, v in params.items()}
    h = torch.tensor(h0, dtype=dtype, requires_grad=True)
    acts: Dict[str, np.ndarray] = {}
    x = h
    if "stem_kernel" in p:
        x, z0, a0 = stem(p, h, stride=2 if params["stem_kernel"].shape[0] == 7 else 4)
        if keep:
            acts["stem_z"], acts["stem_a"], acts["h0"] = z0.detach().numpy(), a0.detach().numpy(), x.detach().numpy()
    for u, kind in enumerate(_plan(units)):
        s = approx_sign(x)
        wq = magnitude_aware_sign(p[f"kernel{u}"])
        sn = s.permute(0, 3, 1, 2)
        if kind == "down":
            z = torch.nn.functional.conv2d(_same_pad_stride2(sn), wq.permute(3, 2, 0, 1), stride=2).permute(0, 2, 3, 1)
            pooled = torch.nn.functional.avg_pool2d(x.permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1)
            t = pooled @ p[f"sc_kernel{u}"]                                   # 1x1 conv, no bias
            shortcut, _, _ = _batchnorm(t, p[f"sc_gamma{u}"], p[f"sc_beta{u}"])
        else:
            z = torch.nn.functional.conv2d(sn, wq.permute(3, 2, 0, 1), padding=1).permute(0, 2, 3, 1)
            shortcut = x
        bn, mean, var = _batchnorm(z, p[f"gamma{u}"], p[f"beta{u}"])
        if keep:
            with torch.no_grad():
                alpha = np.abs(params[f"kernel{u}"].astype(np.float64)).mean(axis=(0, 1, 2))
                acts[f"y{u}"] = np.rint((z.detach().numpy() / alpha)).astype(np.int64)   # integer conv output
                acts[f"alpha{u}"] = alpha
                acts[f"mean_z{u}"], acts[f"var_z{u}"] = mean.detach().numpy(), var.detach().numpy()
        x = bn + shortcut
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
