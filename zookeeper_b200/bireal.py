"""Bi-Real-Net style binary residual units on libzkb200.

The building block BASELINE.json configs[3] / [4] (larq-zoo BinaryAlexNet / Bi-Real-Net-18) add to
the example's BinaryNet: a binary 3x3 conv whose INPUT keeps a float path alive,

    h_next = BatchNormalization()(QuantConv2D(3x3, "same", input_quantizer="approx_sign",
                 kernel_quantizer="magnitude_aware_sign", kernel_constraint="weight_clip")(h)) + h

followed here by the zoo models' head (global average pooling, float Dense, softmax + sparse CE)
and Adam with the latent-weight clip.  The reference reaches such models through the hook the
example uses for BinaryNet -- a ``@factory`` model built from larq layers with the quantizer names in
a ``kwhparams`` dict (/root/reference/examples/larq_experiment.py:41-57, 59-103, compiled and fitted
at :117-122, :141-153); the operators themselves are larq / larq-zoo's (not in the reference tree).

Everything arithmetic is a C-ABI call (``include/zkb200.h``): the binary contraction runs on the
same tcgen05 GEMMs as BinaryNet (int8 forward with exact int32 accumulation, bf16 hi/lo backward),
the per-channel weight magnitude folds into the BatchNorm gain, and only the quantizer backward,
the gain BatchNorm, the shortcut add and the head are new kernels (``csrc/bireal.cu``).  No CPU or
PyTorch fallback: PyTorch holds device memory only.

What is NOT here: the float 7x7/2 stem conv, stride-2 binary convs with their average-pool + 1x1
float shortcut, and a Keras functional-API front end -- the unit takes float activations [B,H,W,C]
as input (``zkb_maxpool3s2_f32_*``, the stem's pool, is available as a kernel).
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import numpy as np
import torch

from . import native as zn

BN_EPS = 1e-3


class _Unit:
    pass


class BiRealUnitsEngine:
    """``units`` residual binary units of ``channels`` channels on a [B, H, W, channels] float map,
    then global average pooling and a float classifier.  ``channels`` must be a multiple of 128 (the
    int8 forward GEMM reads 128 channels per shared-memory row)."""

    def __init__(self, units: int, channels: int, hw, batch_size: int, num_classes: int,
                 learning_rate: float = 5e-3, bn_momentum: float = 0.99, beta_1: float = 0.9,
                 beta_2: float = 0.999, epsilon: float = 1e-7, device: Optional[torch.device] = None):
        zn.load()
        if not torch.cuda.is_available():
            raise RuntimeError("BiRealUnitsEngine needs a CUDA device (B200); there is no CPU fallback")
        if channels % 128 != 0 or channels > 1024:
            raise ValueError(f"channels={channels} must be a multiple of 128 (at most 1024)")
        if units < 1:
            raise ValueError("at least one residual unit")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.U, self.C, self.B, self.NC = int(units), int(channels), int(batch_size), int(num_classes)
        self.H, self.W = int(hw[0]), int(hw[1])
        self.lr, self.b1, self.b2, self.eps = float(learning_rate), beta_1, beta_2, epsilon
        self.momentum = float(bn_momentum)
        self.step_count = 0
        self._allocate()

    # ------------------------------------------------------------------ memory
    def _allocate(self):
        dev, B, H, W, C, U, NC = self.device, self.B, self.H, self.W, self.C, self.U, self.NC
        f32, bf16 = torch.float32, torch.bfloat16
        ksz = 9 * C * C
        self.n_kernel = U * ksz                       # latent binary kernels first: one clipped Adam launch
        n_rest = U * 2 * C + C * NC + NC
        n_rest_pad = (n_rest + 3) // 4 * 4
        n = self.n_kernel + n_rest_pad
        self.params = torch.zeros(n, dtype=f32, device=dev)
        self.grads = torch.zeros(n, dtype=f32, device=dev)
        self.adam_m = torch.zeros(n, dtype=f32, device=dev)
        self.adam_v = torch.zeros(n, dtype=f32, device=dev)
        self.zeros_c = torch.zeros(C, dtype=f32, device=dev)
        self.ones_c = torch.ones(C, dtype=f32, device=dev)
        self.scratch_c = torch.zeros(2 * C, dtype=f32, device=dev)
        self.acc = torch.zeros(U * 4 * C + 2, dtype=torch.float64, device=dev)
        self.correct = torch.zeros(1, dtype=torch.int32, device=dev)
        self.labels = torch.zeros(B, dtype=torch.int32, device=dev)
        self.h = [torch.zeros(B, H, W, C, dtype=f32, device=dev) for _ in range(U + 1)]
        self.dh = [torch.zeros(B, H, W, C, dtype=f32, device=dev) for _ in range(U + 1)]
        self.units = []
        off = self.n_kernel
        for u in range(U):
            X = _Unit()
            X.kernel = self.params[u * ksz:(u + 1) * ksz]
            X.dkernel = self.grads[u * ksz:(u + 1) * ksz]
            X.gamma, X.dgamma = self.params[off:off + C], self.grads[off:off + C]
            X.beta, X.dbeta = self.params[off + C:off + 2 * C], self.grads[off + C:off + 2 * C]
            off += 2 * C
            a = self.acc[u * 4 * C:(u + 1) * 4 * C]
            X.sum, X.sumsq, X.sum_d, X.sum_dx = a[:C], a[C:2 * C], a[2 * C:3 * C], a[3 * C:]
            X.alpha, X.mean_y, X.rs_bwd, X.gain, X.mean_d, X.mean_dx = (torch.zeros(C, dtype=f32, device=dev) for _ in range(6))
            X.mm = torch.zeros(C, dtype=f32, device=dev)
            X.mv = torch.ones(C, dtype=f32, device=dev)
            X.s = torch.zeros(B, H, W, C, dtype=bf16, device=dev)
            X.s8 = torch.zeros(B, H, W, C, dtype=torch.int8, device=dev)
            X.y = torch.zeros(B, H, W, C, dtype=torch.int16, device=dev)
            X.dy_hi = torch.zeros(B, H, W, C, dtype=bf16, device=dev)
            X.dy_lo = torch.zeros(B, H, W, C, dtype=bf16, device=dev)
            X.ds = torch.zeros(B, H, W, C, dtype=f32, device=dev)
            X.w_fwd = torch.zeros(C, 9 * C, dtype=bf16, device=dev)
            X.w_fwd8 = torch.zeros(C, 9 * C, dtype=torch.int8, device=dev)
            X.w_dgrad = torch.zeros(C, 9 * C, dtype=bf16, device=dev)
            self.units.append(X)
        self.dense_w, self.d_dense_w = self.params[off:off + C * NC], self.grads[off:off + C * NC]
        off += C * NC
        self.dense_b, self.d_dense_b = self.params[off:off + NC], self.grads[off:off + NC]
        self.n_rest = n_rest_pad
        self.pooled = torch.zeros(B, C, dtype=f32, device=dev)
        self.dpooled = torch.zeros(B, C, dtype=f32, device=dev)
        self.logits = torch.zeros(B, NC, dtype=f32, device=dev)
        self.dlogits = torch.zeros(B, NC, dtype=f32, device=dev)

    # ------------------------------------------------------------------ parameters in / out
    def load_params(self, p: Dict[str, np.ndarray]):
        """``kernel{u}`` HWIO [3,3,C,C], ``gamma{u}``, ``beta{u}`` [C], ``dense_kernel`` [C,classes],
        ``dense_bias`` [classes]."""
        for u, X in enumerate(self.units):
            X.kernel.copy_(torch.as_tensor(np.ascontiguousarray(p[f"kernel{u}"]), dtype=torch.float32).reshape(-1))
            X.gamma.copy_(torch.as_tensor(p[f"gamma{u}"], dtype=torch.float32))
            X.beta.copy_(torch.as_tensor(p[f"beta{u}"], dtype=torch.float32))
        self.dense_w.copy_(torch.as_tensor(np.ascontiguousarray(p["dense_kernel"]), dtype=torch.float32).reshape(-1))
        self.dense_b.copy_(torch.as_tensor(p["dense_bias"], dtype=torch.float32))

    def _named(self, arena_of) -> Dict[str, np.ndarray]:
        C, NC = self.C, self.NC
        out = {}
        for u, X in enumerate(self.units):
            k, g, b = arena_of(X, "kernel"), arena_of(X, "gamma"), arena_of(X, "beta")
            out[f"kernel{u}"] = k.detach().cpu().numpy().reshape(3, 3, C, C).copy()
            out[f"gamma{u}"] = g.detach().cpu().numpy().copy()
            out[f"beta{u}"] = b.detach().cpu().numpy().copy()
        return out

    def get_params(self) -> Dict[str, np.ndarray]:
        out = self._named(lambda X, k: getattr(X, k))
        out["dense_kernel"] = self.dense_w.detach().cpu().numpy().reshape(self.C, self.NC).copy()
        out["dense_bias"] = self.dense_b.detach().cpu().numpy().copy()
        return out

    def get_grads(self) -> Dict[str, np.ndarray]:
        out = self._named(lambda X, k: getattr(X, "d" + k))
        out["dense_kernel"] = self.d_dense_w.detach().cpu().numpy().reshape(self.C, self.NC).copy()
        out["dense_bias"] = self.d_dense_b.detach().cpu().numpy().copy()
        out["h0"] = self.dh[0].detach().cpu().numpy().copy()
        return out

    # ------------------------------------------------------------------ one step
    def _call(self, name, *args):
        self._n_launch += 1
        zn.call(name, *args, self._stream)

    def prepare_weights(self):
        C = self.C
        for X in self.units:
            self._call("zkb_prep_conv_weights", zn.ptr(X.kernel), 3, C, C, zn.ptr(X.w_fwd), zn.ptr(X.w_dgrad), zn.ptr(X.w_fwd8))
            self._call("zkb_weight_abs_mean", zn.ptr(X.kernel), 9 * C, C, zn.ptr(X.alpha))

    def forward(self, training: bool = True):
        B, H, W, C = self.B, self.H, self.W, self.C
        rows = B * H * W
        for u, X in enumerate(self.units):
            # approx_sign forward = sign(h) (+1 at 0): the BatchNorm+sign kernel with an identity affine
            self._call("zkb_bn_apply_sign", zn.ptr(self.h[u]), 0, rows, C, zn.ptr(self.zeros_c), zn.ptr(self.ones_c),
                       zn.ptr(self.zeros_c), zn.ptr(X.s), 0, zn.ptr(X.s8))
            self._call("zkb_bconv2d_fwd_i8_stats", zn.ptr(X.s8), zn.ptr(X.w_fwd8), zn.ptr(X.y), B, H, W, C, C, 3,
                       zn.ptr(X.sum), zn.ptr(X.sumsq))
            self._call("zkb_bn_gain_finalize", zn.ptr(X.sum), zn.ptr(X.sumsq), float(rows), zn.ptr(X.alpha), zn.ptr(X.gamma),
                       BN_EPS, self.momentum, 1, C, zn.ptr(X.mean_y), zn.ptr(X.rs_bwd), zn.ptr(X.gain),
                       zn.ptr(X.mm) if training else 0, zn.ptr(X.mv) if training else 0)
            self._call("zkb_bn_gain_add", zn.ptr(X.y), rows, C, zn.ptr(X.mean_y), zn.ptr(X.gain), zn.ptr(X.beta),
                       zn.ptr(self.h[u]), zn.ptr(self.h[u + 1]))
        self._call("zkb_global_avgpool_fwd", zn.ptr(self.h[self.U]), B, H * W, C, zn.ptr(self.pooled))
        self._call("zkb_dense_f32_fwd", zn.ptr(self.pooled), zn.ptr(self.dense_w), zn.ptr(self.dense_b), zn.ptr(self.logits),
                   B, C, self.NC, self.NC)

    def backward(self):
        B, H, W, C, NC = self.B, self.H, self.W, self.C, self.NC
        rows = B * H * W
        self._call("zkb_softmax_ce", zn.ptr(self.logits), NC, zn.ptr(self.labels), B, NC, 1.0 / B, zn.ptr(self.dlogits),
                   self.acc[self.U * 4 * C:].data_ptr(), zn.ptr(self.correct))
        self._call("zkb_dense_f32_bwd", zn.ptr(self.pooled), zn.ptr(self.dense_w), zn.ptr(self.dlogits), NC, zn.ptr(self.dpooled),
                   zn.ptr(self.d_dense_w), zn.ptr(self.d_dense_b), B, C, NC)
        self._call("zkb_global_avgpool_bwd", zn.ptr(self.dpooled), B, H * W, C, zn.ptr(self.dh[self.U]))
        for u in range(self.U - 1, -1, -1):
            X = self.units[u]
            d = self.dh[u + 1]
            # BatchNorm (with scale) backward over z = alpha*y: sums with the unmasked gradient
            self._call("zkb_bn_bwd_reduce", zn.ptr(d), zn.ptr(X.y), 1, rows, C, zn.ptr(X.mean_y), zn.ptr(X.rs_bwd), zn.ptr(X.beta), 0,
                       zn.ptr(X.sum_d), zn.ptr(X.sum_dx))
            self._call("zkb_bn_bwd_finalize", zn.ptr(X.sum_d), zn.ptr(X.sum_dx), float(rows), 1.0, C, zn.ptr(X.dbeta),
                       zn.ptr(X.mean_d), zn.ptr(X.mean_dx))
            # dgamma = sum d*xhat: the same finalize with the two sums swapped (its means are scratch)
            self._call("zkb_bn_bwd_finalize", zn.ptr(X.sum_dx), zn.ptr(X.sum_d), float(rows), 1.0, C, zn.ptr(X.dgamma),
                       zn.ptr(self.scratch_c), self.scratch_c[C:].data_ptr())
            self._call("zkb_bn_gain_bwd_apply", zn.ptr(d), zn.ptr(X.y), rows, C, zn.ptr(X.mean_y), zn.ptr(X.rs_bwd), zn.ptr(X.gain),
                       zn.ptr(X.mean_d), zn.ptr(X.mean_dx), zn.ptr(X.dy_hi), zn.ptr(X.dy_lo))
            self._call("zkb_bconv2d_dgrad", zn.ptr(X.dy_hi), zn.ptr(X.dy_lo), zn.ptr(X.w_dgrad), zn.ptr(X.ds), B, H, W, C, C, 3)
            # dy = alpha*dz, so this IS the latent-kernel gradient alpha * wgrad(sign(h), dz); the
            # |w| <= 1 mask of magnitude_aware_sign's ste is applied by the Adam kernel
            self._call("zkb_bconv2d_wgrad", zn.ptr(X.s), zn.ptr(X.dy_hi), zn.ptr(X.dy_lo), zn.ptr(X.dkernel), B, H, W, C, C, 3)
            # approx_sign backward, plus the gradient that arrives over the float shortcut
            self._call("zkb_approx_sign_bwd_add", zn.ptr(X.ds), zn.ptr(self.h[u]), zn.ptr(d), zn.ptr(self.dh[u]), rows * C)

    def _lr_t(self, t: int) -> float:
        f = np.float32
        return float(f(self.lr) * np.sqrt(f(1) - f(self.b2) ** f(t), dtype=f) / (f(1) - f(self.b1) ** f(t)))

    def optimizer_step(self):
        self.step_count += 1
        lr_t = self._lr_t(self.step_count)
        nk = self.n_kernel
        self._call("zkb_adam_clip", zn.ptr(self.params), zn.ptr(self.grads), zn.ptr(self.adam_m), zn.ptr(self.adam_v),
                   nk, lr_t, self.b1, self.b2, self.eps, 1.0, 1, 0)
        self._call("zkb_adam_clip", self.params[nk:].data_ptr(), self.grads[nk:].data_ptr(), self.adam_m[nk:].data_ptr(),
                   self.adam_v[nk:].data_ptr(), self.n_rest, lr_t, self.b1, self.b2, self.eps, 1.0, 0, 0)

    def train_step(self, h0: torch.Tensor, labels: torch.Tensor, apply_update: bool = True):
        """h0: float32 [B,H,W,C] activations entering the first unit; labels: integer [B]."""
        self._stream = torch.cuda.current_stream(self.device).cuda_stream
        self._n_launch = 0
        self.acc.zero_()
        self.correct.zero_()
        self.grads.zero_()
        self.h[0].copy_(h0)
        self.labels.copy_(labels.to(torch.int32))
        self.prepare_weights()
        self.forward(training=True)
        self.backward()
        if apply_update:
            self.optimizer_step()
        self.launches_per_step = self._n_launch

    def last_loss(self) -> float:
        return float(self.acc[self.U * 4 * self.C].item()) / self.B

    def last_accuracy(self) -> float:
        return float(self.correct.item()) / self.B
