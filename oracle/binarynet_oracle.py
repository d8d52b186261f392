"""CPU oracle for the BinaryNet train hot path -- TEST INFRASTRUCTURE ONLY.

This file restates, in plain numpy / torch-CPU, the arithmetic that the reference delegates to
TensorFlow + Larq for ``examples/larq_experiment.py`` (reference file:line cited per function).
It is imported only by ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline
legs; the product path (``zookeeper_b200``) never imports it and has no CPU fallback.

PARITY UNPINNED: TensorFlow, Keras and Larq are third-party dependencies that are not vendored in
/root/reference, are not installed in this image and cannot be installed (no network); the
reference holds no golden vector, known-answer test or fixture for any numeric quantity of this
path (zookeeper/tf has no tests; CI only runs ``epochs=0``, .github/workflows/unittest.yml:41-42).
The operator semantics below follow the published behaviour of
  - larq ``ste_sign`` / ``QuantConv2D`` / ``QuantDense`` / ``weight_clip`` (larq is unpinned in
    the reference: .github/workflows/unittest.yml:37),
  - ``tf.keras`` ``BatchNormalization(scale=False)``, ``MaxPool2D``, ``Adam``, softmax +
    sparse categorical cross-entropy (TensorFlow 1.14 .. 2.3, unittest.yml:14),
as summarised in SURVEY.md Appendix A.  The golden vectors under tests/golden/ are produced by
THIS file (oracle/make_golden.py) and pin the oracle against drift, not against TensorFlow.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

BN_EPS = 1e-3        # keras BatchNormalization default epsilon
BN_MOMENTUM = 0.99   # keras BatchNormalization default momentum
ADAM_B1, ADAM_B2, ADAM_EPS = 0.9, 0.999, 1e-7  # tf.keras.optimizers.Adam defaults


# --------------------------------------------------------------------------------------------
# Preprocessing: examples/larq_experiment.py:24-37 (PadCropAndFlip.input / .output)
# --------------------------------------------------------------------------------------------
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _mix64(z: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser on uint64 arrays (wrap-around arithmetic)."""
    with np.errstate(over="ignore"):
        z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def augmentation_params(seed: int, step: int, batch: int, pad_size: int, out_h: int, out_w: int) -> np.ndarray:
    """Per-sample (ox, oy, flip) of the counter-based generator the CUDA kernel uses.

    tf.image.random_crop / random_flip_left_right (larq_experiment.py:30-31) draw from TF's RNG,
    which cannot be reproduced; the distribution is the same (uniform offsets, p = 1/2 flip).
    """
    with np.errstate(over="ignore"):
        key = _mix64(np.asarray([np.uint64(seed) ^ (np.uint64(step) * np.uint64(0xD1342543DE82EF95))], dtype=np.uint64))
        r = _mix64(key + np.arange(batch, dtype=np.uint64))
    ox = (r & np.uint64(0xFFFF)) % np.uint64(pad_size - out_w + 1)
    oy = ((r >> np.uint64(16)) & np.uint64(0xFFFF)) % np.uint64(pad_size - out_h + 1)
    flip = (r >> np.uint64(32)) & np.uint64(1)
    return np.stack([ox, oy, flip], axis=1).astype(np.int8)


def _crop_or_pad_offset(src: int, dst: int) -> int:
    # tf.image.resize_with_crop_or_pad centres: pad (dst-src)//2 before, or crop (src-dst)//2.
    return (dst - src) // 2 if dst >= src else -((src - dst) // 2)


def preprocess(images_u8: np.ndarray, out_hw: Tuple[int, int], training: bool, pad_size: int = 0,
               aug: Optional[np.ndarray] = None, binarize: bool = False) -> np.ndarray:
    """uint8 NHWC -> float32 NHWC in [-1, 1]  (larq_experiment.py:24-34).

    training: centre zero-pad to pad_size, crop out_hw at (oy, ox), optional left-right flip.
    eval    : centre crop/pad to out_hw.  Then float32(x) / 127.5 - 1.0 (true division).
    """
    n, ih, iw, c = images_u8.shape
    oh, ow = out_hw
    out = np.zeros((n, oh, ow, c), dtype=np.uint8)
    for b in range(n):
        if training:
            canvas = np.zeros((pad_size, pad_size, c), dtype=np.uint8)
            off_h, off_w = _crop_or_pad_offset(ih, pad_size), _crop_or_pad_offset(iw, pad_size)
        else:
            canvas = np.zeros((oh, ow, c), dtype=np.uint8)
            off_h, off_w = _crop_or_pad_offset(ih, oh), _crop_or_pad_offset(iw, ow)
        ch, cw = canvas.shape[:2]
        # paste the source image at (off_h, off_w) (negative offsets crop)
        sh0, sw0 = max(0, -off_h), max(0, -off_w)
        dh0, dw0 = max(0, off_h), max(0, off_w)
        hh, ww = min(ih - sh0, ch - dh0), min(iw - sw0, cw - dw0)
        canvas[dh0:dh0 + hh, dw0:dw0 + ww] = images_u8[b, sh0:sh0 + hh, sw0:sw0 + ww]
        if training:
            ox, oy, flip = (int(v) for v in aug[b])
            img = canvas[oy:oy + oh, ox:ox + ow]
            if flip:
                img = img[:, ::-1]
        else:
            img = canvas
        out[b] = img
    x = out.astype(np.float32) / np.float32(127.5) - np.float32(1.0)
    if binarize:
        x = np.where(x >= 0, np.float32(1.0), np.float32(-1.0))
    return x


# --------------------------------------------------------------------------------------------
# larq quantizer ste_sign (kwhparams, larq_experiment.py:52-57): sign(0) = +1, grad = dy*[|x|<=1]
# --------------------------------------------------------------------------------------------
class _SteSign(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(x)
        return torch.where(x >= 0, torch.ones_like(x), -torch.ones_like(x))

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        return dy * (x.abs() <= 1).to(dy.dtype)


def ste_sign(x: torch.Tensor) -> torch.Tensor:
    return _SteSign.apply(x)


# --------------------------------------------------------------------------------------------
# MaxPool2D(2,2) 'valid' (larq_experiment.py:74,83,92): first maximum in row-major window order
# receives the gradient (TF CPU kernel replaces the running max on strict '>').
# --------------------------------------------------------------------------------------------
def _pool_windows(x: torch.Tensor) -> torch.Tensor:
    n, h, w, c = x.shape
    ph, pw = h // 2, w // 2
    x = x[:, : 2 * ph, : 2 * pw]
    x = x.reshape(n, ph, 2, pw, 2, c).permute(0, 1, 3, 2, 4, 5).reshape(n, ph, pw, 4, c)
    return x


def pool_argmax_first(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    win = _pool_windows(x)
    best = win[:, :, :, 0]
    arg = torch.zeros_like(best, dtype=torch.int64)
    for k in range(1, 4):
        better = win[:, :, :, k] > best
        best = torch.where(better, win[:, :, :, k], best)
        arg = torch.where(better, torch.full_like(arg, k), arg)
    return best, arg


class _MaxPoolFirst(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        best, arg = pool_argmax_first(x)
        ctx.save_for_backward(arg)
        ctx.in_shape = x.shape
        return best

    @staticmethod
    def backward(ctx, dy):
        (arg,) = ctx.saved_tensors
        n, h, w, c = ctx.in_shape
        ph, pw = h // 2, w // 2
        onehot = torch.nn.functional.one_hot(arg, 4).permute(0, 1, 2, 4, 3).to(dy.dtype)  # n,ph,pw,4,c
        dwin = onehot * dy.unsqueeze(3)
        dx = torch.zeros(n, h, w, c, dtype=dy.dtype)
        dwin = dwin.reshape(n, ph, pw, 2, 2, c).permute(0, 1, 3, 2, 4, 5).reshape(n, 2 * ph, 2 * pw, c)
        dx[:, : 2 * ph, : 2 * pw] = dwin
        return dx


def maxpool2(x: torch.Tensor) -> torch.Tensor:
    return _MaxPoolFirst.apply(x)


# --------------------------------------------------------------------------------------------
# Model description: BinaryNet.build, examples/larq_experiment.py:51-103
# --------------------------------------------------------------------------------------------
@dataclass
class LayerSpec:
    kind: str            # "conv" | "dense"
    cin: int
    cout: int
    ksize: int = 3
    padding: str = "same"
    quant_input: bool = True
    pool: bool = False


def binarynet_layers(input_shape=(28, 28, 1), filters=128, dense_units=1024, num_classes=10,
                     ksize=3) -> List[LayerSpec]:
    h, w, c = input_shape
    f = filters
    layers = [
        LayerSpec("conv", c, f, ksize, "valid", quant_input=False),          # :62-69
        LayerSpec("conv", f, f, ksize, "same", pool=True),                    # :71-74
        LayerSpec("conv", f, 2 * f, ksize, "same"),                           # :76-78
        LayerSpec("conv", 2 * f, 2 * f, ksize, "same", pool=True),            # :80-83
        LayerSpec("conv", 2 * f, 4 * f, ksize, "same"),                       # :85-87
        LayerSpec("conv", 4 * f, 4 * f, ksize, "same", pool=True),            # :89-92
    ]
    h, w = h - ksize + 1, w - ksize + 1
    for _ in range(3):
        h, w = h // 2, w // 2
    flat = h * w * 4 * f                                                       # Flatten :94
    layers += [
        LayerSpec("dense", flat, dense_units),                                 # :95
        LayerSpec("dense", dense_units, dense_units),                          # :97
        LayerSpec("dense", dense_units, num_classes),                          # :99
    ]
    return layers


def glorot_uniform_params(layers: Sequence[LayerSpec], seed: int = 0) -> Dict[str, np.ndarray]:
    """Keras default kernel initialiser, HWIO / [K, N]; beta = 0 (BatchNormalization center)."""
    rng = np.random.default_rng(seed)
    params = {}
    for i, l in enumerate(layers):
        if l.kind == "conv":
            shape = (l.ksize, l.ksize, l.cin, l.cout)
            fan_in, fan_out = l.ksize * l.ksize * l.cin, l.ksize * l.ksize * l.cout
        else:
            shape = (l.cin, l.cout)
            fan_in, fan_out = l.cin, l.cout
        lim = np.sqrt(6.0 / (fan_in + fan_out))
        params[f"kernel{i}"] = rng.uniform(-lim, lim, size=shape).astype(np.float32)
        params[f"beta{i}"] = np.zeros(l.cout, dtype=np.float32)
    return params


def init_moving_stats(layers: Sequence[LayerSpec]) -> Dict[str, np.ndarray]:
    st = {}
    for i, l in enumerate(layers):
        st[f"moving_mean{i}"] = np.zeros(l.cout, dtype=np.float32)
        st[f"moving_var{i}"] = np.ones(l.cout, dtype=np.float32)
    return st


def _conv_nhwc(x: torch.Tensor, w_hwio: torch.Tensor, padding: str) -> torch.Tensor:
    pad = w_hwio.shape[0] // 2 if padding == "same" else 0
    y = torch.nn.functional.conv2d(x.permute(0, 3, 1, 2), w_hwio.permute(3, 2, 0, 1), padding=pad)
    return y.permute(0, 2, 3, 1)


@dataclass
class StepResult:
    loss: float
    accuracy: float
    grads: Dict[str, np.ndarray]
    logits: np.ndarray
    acts: Dict[str, np.ndarray] = field(default_factory=dict)     # per-layer intermediates
    batch_stats: Dict[str, np.ndarray] = field(default_factory=dict)


def forward_backward(params: Dict[str, np.ndarray], layers: Sequence[LayerSpec], x: np.ndarray,
                     labels: np.ndarray, dtype=torch.float64, keep_acts: bool = False) -> StepResult:
    """One training-mode forward + backward of BinaryNet (larq_experiment.py:59-103, 117).

    QuantConv2D/QuantDense: y = op(ste_sign(x) if input_quantizer else x, ste_sign(W)), zero "same"
    padding after quantisation; BatchNormalization(scale=False) with batch statistics (biased
    variance); loss = mean_b( -log softmax(z_b)[label_b] ).
    """
    p = {k: torch.tensor(v, dtype=dtype, requires_grad=True) for k, v in params.items()}
    h = torch.tensor(x, dtype=dtype)
    acts, stats = {}, {}
    flat_done = False
    for i, l in enumerate(layers):
        if l.kind == "dense" and not flat_done:
            h = h.reshape(h.shape[0], -1)          # Flatten, (h, w, c) order
            flat_done = True
        a_in = ste_sign(h) if l.quant_input else h
        wq = ste_sign(p[f"kernel{i}"])
        y = _conv_nhwc(a_in, wq, l.padding) if l.kind == "conv" else a_in @ wq
        if keep_acts:
            acts[f"y{i}"] = y.detach().numpy().copy()
        if l.pool:
            y = maxpool2(y)
            if keep_acts:
                acts[f"p{i}"] = y.detach().numpy().copy()
        red = tuple(range(y.dim() - 1))
        mean = y.mean(dim=red)
        var = ((y - mean) ** 2).mean(dim=red)
        stats[f"mean{i}"] = mean.detach().numpy().copy()
        stats[f"var{i}"] = var.detach().numpy().copy()
        stats[f"count{i}"] = np.asarray(float(np.prod(y.shape[:-1])))
        h = (y - mean) * torch.rsqrt(var + BN_EPS) + p[f"beta{i}"]
        if keep_acts:
            acts[f"a{i}"] = h.detach().numpy().copy()
    logits = h
    lab = torch.tensor(labels, dtype=torch.int64)
    loss = torch.nn.functional.cross_entropy(logits, lab, reduction="mean")
    loss.backward()
    acc = float((logits.argmax(dim=1) == lab).double().mean())
    grads = {k: v.grad.detach().numpy().copy() for k, v in p.items()}
    return StepResult(float(loss.detach()), acc, grads, logits.detach().numpy().copy(), acts, stats)


def update_moving_stats(moving: Dict[str, np.ndarray], layers: Sequence[LayerSpec],
                        batch_stats: Dict[str, np.ndarray]) -> Dict[str, np.ndarray]:
    """moving <- moving*0.99 + batch*0.01.  4-D (conv) layers run TF's fused kernel, which feeds the
    Bessel-corrected variance to the moving average; 2-D (dense) layers use the biased one
    (SURVEY.md A.4)."""
    out = {}
    for i, l in enumerate(layers):
        mean, var, n = batch_stats[f"mean{i}"], batch_stats[f"var{i}"], float(batch_stats[f"count{i}"])
        if l.kind == "conv" and n > 1:
            var = var * (n / (n - 1.0))
        out[f"moving_mean{i}"] = (moving[f"moving_mean{i}"] * BN_MOMENTUM + mean * (1 - BN_MOMENTUM)).astype(np.float32)
        out[f"moving_var{i}"] = (moving[f"moving_var{i}"] * BN_MOMENTUM + var * (1 - BN_MOMENTUM)).astype(np.float32)
    return out


def adam_clip_step(params, grads, m, v, step: int, lr: float = 5e-3, dtype=np.float64):
    """tf.keras Adam(lr) (larq_experiment.py:120-122) then larq weight_clip on kernels (:55,66).

    m <- b1 m + (1-b1) g ; v <- b2 v + (1-b2) g^2 ; w <- w - lr_t m / (sqrt(v) + eps) with
    lr_t = lr sqrt(1-b2^t)/(1-b1^t) ; kernels are then clipped to [-1, 1].
    With dtype=np.float32 every constant and operation is rounded to fp32 in the order of TF's
    ApplyAdam CPU functor (alpha = lr*sqrt(1-b2^t)/(1-b1^t); m += (g-m)(1-b1); v += (g*g-v)(1-b2);
    var -= m*alpha/(sqrt(v)+eps)); with np.float64 it is the exact-arithmetic reference.
    """
    f = dtype
    b1, b2, eps = f(ADAM_B1), f(ADAM_B2), f(ADAM_EPS)
    lr_t = f(lr) * np.sqrt(f(1) - b2 ** f(step), dtype=f) / (f(1) - b1 ** f(step))
    new_p, new_m, new_v = {}, {}, {}
    for k in params:
        g = grads[k].astype(dtype)
        w = params[k].astype(dtype)
        if k.startswith("kernel"):
            g = g * (np.abs(w) <= 1)          # ste_sign kernel quantizer mask (always 1 after clip)
        mm = m[k].astype(dtype)
        vv = v[k].astype(dtype)
        mm = mm + (g - mm) * (f(1) - b1)
        vv = vv + (g * g - vv) * (f(1) - b2)
        w = w - (mm * lr_t) / (np.sqrt(vv) + eps)
        if k.startswith("kernel"):
            w = np.clip(w, -1.0, 1.0)
        new_p[k], new_m[k], new_v[k] = w.astype(np.float32), mm.astype(np.float32), vv.astype(np.float32)
    return new_p, new_m, new_v


def predict(params, moving, layers, x: np.ndarray, dtype=torch.float64) -> np.ndarray:
    """Inference-mode forward (BatchNorm with moving statistics), larq_experiment.py:133-139."""
    h = torch.tensor(x, dtype=dtype)
    flat_done = False
    with torch.no_grad():
        for i, l in enumerate(layers):
            if l.kind == "dense" and not flat_done:
                h = h.reshape(h.shape[0], -1)
                flat_done = True
            a_in = ste_sign(h) if l.quant_input else h
            wq = ste_sign(torch.tensor(params[f"kernel{i}"], dtype=dtype))
            y = _conv_nhwc(a_in, wq, l.padding) if l.kind == "conv" else a_in @ wq
            if l.pool:
                y = pool_argmax_first(y)[0]
            mm = torch.tensor(moving[f"moving_mean{i}"], dtype=dtype)
            mv = torch.tensor(moving[f"moving_var{i}"], dtype=dtype)
            h = (y - mm) * torch.rsqrt(mv + BN_EPS) + torch.tensor(params[f"beta{i}"], dtype=dtype)
    return h.numpy()


# --------------------------------------------------------------------------------------------
# Bit-level references (north_star items 2 and 3): sign + pack along K, XNOR-popcount GEMM.
# --------------------------------------------------------------------------------------------
def sign_pack(x: np.ndarray) -> np.ndarray:
    """[rows, K] float -> [rows, K/32] uint32; bit j of word w is 1 iff x[row, 32 w + j] >= 0."""
    rows, k = x.shape
    assert k % 32 == 0
    bits = (x >= 0).astype(np.uint32).reshape(rows, k // 32, 32)
    weights = (np.uint32(1) << np.arange(32, dtype=np.uint32))
    return (bits * weights).sum(axis=2, dtype=np.uint64).astype(np.uint32)


def xnor_popcount_gemm(a_packed: np.ndarray, b_packed: np.ndarray, k: int) -> np.ndarray:
    """y[m, n] = k - 2 * popcount(a[m] xor b[n]) == sum_k (+-1)(+-1), in pure integers."""
    x = a_packed[:, None, :] ^ b_packed[None, :, :]
    pop = np.unpackbits(x.view(np.uint8), axis=-1).sum(axis=-1, dtype=np.int64)
    return (k - 2 * pop).astype(np.int32)
