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

A unit is either "plain" (C -> C, stride 1, identity shortcut) or "down" (C -> 2C, stride 2 with
TensorFlow 'same' padding; shortcut = AveragePooling2D(2, 2) -> Conv2D 1x1 without bias ->
BatchNormalization, all float), so the stack below is the body of a Bi-Real-Net from any stage on.

``BiRealNetEngine`` puts the zoo model's stem in front of such a stack -- float Conv2D(64, 7,
strides=2, "same") on the image, BatchNormalization, MaxPool2D(3, 2, "same") -- which makes the whole
Bi-Real-Net-18 (``bireal18_plan()``: four 64-channel units, then three stages that open with a
down-sampling unit) one engine.  Channel counts are multiples of 64; layers whose input width is
not a multiple of 128 (the 64-channel stage) run the bf16 +-1 forward GEMM instead of the int8 one
(same integers, half the tensor-pipe rate).

The Keras-shaped handle (``compile`` / ``summary`` / ``fit``) is ``zookeeper_b200.tf.zoo.BiRealNet18``.
What is NOT here: an inference path.  The 1x1 float conv of the down-sampling shortcut and the stem
conv run on register-tiled FP32 FMA kernels, not on the tensor cores.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import numpy as np
import torch

from . import native as zn
from .dataparallel import DataParallel

BN_EPS = 1e-3


class _Unit:
    pass


class BiRealUnitsEngine:
    """A stack of binary residual units on a [B, H, W, channels] float map, then global average
    pooling and a float classifier.  ``units``: a count of plain units, or a sequence of "plain" /
    "down".  Channel counts must be multiples of 64 (a shared-memory row of the bf16 GEMM operands;
    multiples of 128 take the int8 forward); "down" needs an even map."""

    def __init__(self, units, channels: int, hw, batch_size: int, num_classes: int,
                 learning_rate: float = 5e-3, bn_momentum: float = 0.99, beta_1: float = 0.9,
                 beta_2: float = 0.999, epsilon: float = 1e-7, device: Optional[torch.device] = None,
                 data_parallel: Optional[DataParallel] = None):
        zn.load()
        if not torch.cuda.is_available():
            raise RuntimeError("BiRealUnitsEngine needs a CUDA device (B200); there is no CPU fallback")
        # one process per GPU: the batch shards by sample, BatchNorm statistics stay per replica, the
        # flat gradient arena (44 MB for Bi-Real-Net-18) is summed over the replicas after the backward
        self.dp = data_parallel if data_parallel is not None else DataParallel()
        self.world = self.dp.world
        self.plan = ["plain"] * units if isinstance(units, int) else [str(u) for u in units]
        if not self.plan or any(u not in ("plain", "down") for u in self.plan):
            raise ValueError('units: a positive count or a sequence of "plain" / "down"')
        if channels % 64 != 0:
            raise ValueError(f"channels={channels} must be a multiple of 64")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.U, self.C, self.B, self.NC = len(self.plan), int(channels), int(batch_size), int(num_classes)
        self.H, self.W = int(hw[0]), int(hw[1])
        self.lr, self.b1, self.b2, self.eps = float(learning_rate), beta_1, beta_2, epsilon
        self.momentum = float(bn_momentum)
        self.step_count = 0
        self._allocate()

    # ------------------------------------------------------------------ memory
    def _allocate(self):
        dev, B, NC = self.device, self.B, self.NC
        f32, bf16 = torch.float32, torch.bfloat16
        # shapes through the stack
        shapes, h, w, c = [], self.H, self.W, self.C
        for kind in self.plan:
            if kind == "down":
                if h % 2 or w % 2:
                    raise ValueError(f"a down-sampling unit needs an even map, got {h}x{w}")
                shapes.append((h, w, c, h // 2, w // 2, 2 * c))
                h, w, c = h // 2, w // 2, 2 * c
            else:
                shapes.append((h, w, c, h, w, c))
            if c > 1024:
                raise ValueError("at most 1024 channels")
        self.out_hw, self.out_c = (h, w), c
        self.n_kernel = sum(9 * s[2] * s[5] for s in shapes)      # latent binary kernels first: one clipped Adam launch
        n_rest = sum(2 * s[5] + ((s[2] * s[5] + 2 * s[5]) if k == "down" else 0) for s, k in zip(shapes, self.plan)) + c * NC + NC
        n_rest_pad = (n_rest + 3) // 4 * 4
        n = self.n_kernel + n_rest_pad
        self.params = torch.zeros(n, dtype=f32, device=dev)
        self.grads = torch.zeros(n, dtype=f32, device=dev)
        self.adam_m = torch.zeros(n, dtype=f32, device=dev)
        self.adam_v = torch.zeros(n, dtype=f32, device=dev)
        cmax = max(s[5] for s in shapes)
        self.zeros_c = torch.zeros(cmax, dtype=f32, device=dev)
        self.ones_c = torch.ones(cmax, dtype=f32, device=dev)
        self.scratch_c = torch.zeros(2 * cmax, dtype=f32, device=dev)
        n_acc = sum((8 if k == "down" else 4) * s[5] for s, k in zip(shapes, self.plan))
        self.acc = torch.zeros(n_acc + 2, dtype=torch.float64, device=dev)
        self.loss_off = n_acc
        self.correct = torch.zeros(1, dtype=torch.int32, device=dev)
        self.labels = torch.zeros(B, dtype=torch.int32, device=dev)
        self.h = [torch.zeros(B, self.H, self.W, self.C, dtype=f32, device=dev)]
        self.dh = [torch.zeros(B, self.H, self.W, self.C, dtype=f32, device=dev)]
        self.units = []
        koff, off, aoff = 0, self.n_kernel, 0

        def take(nel):
            nonlocal off
            pair = (self.params[off:off + nel], self.grads[off:off + nel])
            off += nel
            return pair

        def vec(nel):
            return torch.zeros(nel, dtype=f32, device=dev)

        for kind, (hi, wi, ci, ho, wo, co) in zip(self.plan, shapes):
            X = _Unit()
            X.kind, X.in_shape, X.out_shape = kind, (hi, wi, ci), (ho, wo, co)
            ksz = 9 * ci * co
            X.kernel, X.dkernel = self.params[koff:koff + ksz], self.grads[koff:koff + ksz]
            koff += ksz
            (X.gamma, X.dgamma), (X.beta, X.dbeta) = take(co), take(co)
            a = self.acc[aoff:aoff + 4 * co]
            aoff += 4 * co
            X.sum, X.sumsq, X.sum_d, X.sum_dx = a[:co], a[co:2 * co], a[2 * co:3 * co], a[3 * co:]
            X.alpha, X.mean_y, X.rs_bwd, X.gain, X.mean_d, X.mean_dx = (vec(co) for _ in range(6))
            X.mm, X.mv = vec(co), torch.ones(co, dtype=f32, device=dev)
            X.s = torch.zeros(B, hi, wi, ci, dtype=bf16, device=dev)
            X.s8 = torch.zeros(B, hi, wi, ci, dtype=torch.int8, device=dev)
            X.y = torch.zeros(B, ho, wo, co, dtype=torch.int16, device=dev)
            X.dy_hi = torch.zeros(B, ho, wo, co, dtype=bf16, device=dev)
            X.dy_lo = torch.zeros(B, ho, wo, co, dtype=bf16, device=dev)
            X.ds = torch.zeros(B, hi, wi, ci, dtype=f32, device=dev)
            X.w_fwd = torch.zeros(co, 9 * ci, dtype=bf16, device=dev)
            X.w_fwd8 = torch.zeros(co, 9 * ci, dtype=torch.int8, device=dev)
            X.w_dgrad = torch.zeros(ci, 9 * co, dtype=bf16, device=dev)
            if kind == "down":
                (X.sc_kernel, X.d_sc_kernel), (X.sc_gamma, X.d_sc_gamma), (X.sc_beta, X.d_sc_beta) = take(ci * co), take(co), take(co)
                a = self.acc[aoff:aoff + 4 * co]
                aoff += 4 * co
                X.sc_sum, X.sc_sumsq, X.sc_sum_d, X.sc_sum_dx = a[:co], a[co:2 * co], a[2 * co:3 * co], a[3 * co:]
                X.sc_mean, X.sc_rs, X.sc_gain, X.sc_mean_d, X.sc_mean_dx = (vec(co) for _ in range(5))
                X.sc_mm, X.sc_mv = vec(co), torch.ones(co, dtype=f32, device=dev)
                X.pool = torch.zeros(B, ho, wo, ci, dtype=f32, device=dev)      # avg-pooled input
                X.dpool = torch.zeros(B, ho, wo, ci, dtype=f32, device=dev)
                X.t = torch.zeros(B, ho, wo, co, dtype=f32, device=dev)         # 1x1 conv output
                X.dt = torch.zeros(B, ho, wo, co, dtype=f32, device=dev)
                X.sc = torch.zeros(B, ho, wo, co, dtype=f32, device=dev)        # BatchNorm(t): the shortcut
                X.dh_sc = torch.zeros(B, hi, wi, ci, dtype=f32, device=dev)
                X.dil_hi = torch.zeros(B, hi, wi, co, dtype=bf16, device=dev)   # zero-inserted dy
                X.dil_lo = torch.zeros(B, hi, wi, co, dtype=bf16, device=dev)
            self.units.append(X)
            self.h.append(torch.zeros(B, ho, wo, co, dtype=f32, device=dev))
            self.dh.append(torch.zeros(B, ho, wo, co, dtype=f32, device=dev))
        (self.dense_w, self.d_dense_w), (self.dense_b, self.d_dense_b) = take(c * NC), take(NC)
        self.n_rest = n_rest_pad
        self.pooled = torch.zeros(B, c, dtype=f32, device=dev)
        self.dpooled = torch.zeros(B, c, dtype=f32, device=dev)
        self.logits = torch.zeros(B, NC, dtype=f32, device=dev)
        self.dlogits = torch.zeros(B, NC, dtype=f32, device=dev)

    # ------------------------------------------------------------------ parameters in / out
    _NAMES = (("kernel", "kernel"), ("gamma", "gamma"), ("beta", "beta"))
    _DOWN_NAMES = (("sc_kernel", "sc_kernel"), ("sc_gamma", "sc_gamma"), ("sc_beta", "sc_beta"))

    def _shape_of(self, X, name):
        (hi, wi, ci), (ho, wo, co) = X.in_shape, X.out_shape
        return {"kernel": (3, 3, ci, co), "sc_kernel": (ci, co)}.get(name, (co,))

    def load_params(self, p: Dict[str, np.ndarray]):
        """``kernel{u}`` HWIO [3,3,Cin,Cout], ``gamma{u}``, ``beta{u}`` [Cout]; down-sampling units also
        ``sc_kernel{u}`` [Cin,Cout], ``sc_gamma{u}``, ``sc_beta{u}``; ``dense_kernel`` [C,classes],
        ``dense_bias`` [classes]."""
        for u, X in enumerate(self.units):
            for key, attr in self._NAMES + (self._DOWN_NAMES if X.kind == "down" else ()):
                getattr(X, attr).copy_(torch.as_tensor(np.ascontiguousarray(p[f"{key}{u}"]), dtype=torch.float32).reshape(-1))
        self.dense_w.copy_(torch.as_tensor(np.ascontiguousarray(p["dense_kernel"]), dtype=torch.float32).reshape(-1))
        self.dense_b.copy_(torch.as_tensor(p["dense_bias"], dtype=torch.float32))

    def _named(self, prefix: str) -> Dict[str, np.ndarray]:
        out = {}
        for u, X in enumerate(self.units):
            for key, attr in self._NAMES + (self._DOWN_NAMES if X.kind == "down" else ()):
                name = ("d_" if prefix and attr.startswith("sc_") else prefix) + attr
                out[f"{key}{u}"] = getattr(X, name).detach().cpu().numpy().reshape(self._shape_of(X, key)).copy()
        return out

    def get_params(self) -> Dict[str, np.ndarray]:
        out = self._named("")
        out["dense_kernel"] = self.dense_w.detach().cpu().numpy().reshape(self.out_c, self.NC).copy()
        out["dense_bias"] = self.dense_b.detach().cpu().numpy().copy()
        return out

    def get_grads(self) -> Dict[str, np.ndarray]:
        out = self._named("d")
        out["dense_kernel"] = self.d_dense_w.detach().cpu().numpy().reshape(self.out_c, self.NC).copy()
        out["dense_bias"] = self.d_dense_b.detach().cpu().numpy().copy()
        out["h0"] = self.dh[0].detach().cpu().numpy().copy()
        return out

    # ------------------------------------------------------------------ one step
    def _call(self, name, *args):
        self._n_launch += 1
        zn.call(name, *args, self._stream)

    def prepare_weights(self):
        for X in self.units:
            ci, co = X.in_shape[2], X.out_shape[2]
            self._call("zkb_prep_conv_weights", zn.ptr(X.kernel), 3, ci, co, zn.ptr(X.w_fwd), zn.ptr(X.w_dgrad), zn.ptr(X.w_fwd8))
            self._call("zkb_weight_abs_mean", zn.ptr(X.kernel), 9 * ci, co, zn.ptr(X.alpha))

    def forward(self, training: bool = True):
        B = self.B
        for u, X in enumerate(self.units):
            (hi, wi, ci), (ho, wo, co) = X.in_shape, X.out_shape
            rows_in, rows = B * hi * wi, B * ho * wo
            mm = (lambda t: zn.ptr(t)) if training else (lambda t: 0)
            # approx_sign forward = sign(h) (+1 at 0): the BatchNorm+sign kernel with an identity affine
            self._call("zkb_bn_apply_sign", zn.ptr(self.h[u]), 0, rows_in, ci, zn.ptr(self.zeros_c), zn.ptr(self.ones_c),
                       zn.ptr(self.zeros_c), zn.ptr(X.s), 0, zn.ptr(X.s8))
            i8 = ci % 128 == 0        # the int8 GEMM reads 128 channels per shared-memory row
            if X.kind == "plain":
                if i8:
                    self._call("zkb_bconv2d_fwd_i8_stats", zn.ptr(X.s8), zn.ptr(X.w_fwd8), zn.ptr(X.y), B, hi, wi, ci, co, 3,
                               zn.ptr(X.sum), zn.ptr(X.sumsq))
                else:
                    self._call("zkb_bconv2d_fwd", zn.ptr(X.s), zn.ptr(X.w_fwd), zn.ptr(X.y), B, hi, wi, ci, co, 3)
                    self._call("zkb_bn_stats", zn.ptr(X.y), 1, rows, co, zn.ptr(X.sum), zn.ptr(X.sumsq))
                shortcut = self.h[u]
            else:
                if i8:
                    self._call("zkb_bconv2d_fwd_i8_s2", zn.ptr(X.s8), zn.ptr(X.w_fwd8), zn.ptr(X.y), B, hi, wi, ci, co, 3)
                else:
                    self._call("zkb_bconv2d_fwd_s2", zn.ptr(X.s), zn.ptr(X.w_fwd), zn.ptr(X.y), B, hi, wi, ci, co, 3)
                self._call("zkb_bn_stats", zn.ptr(X.y), 1, rows, co, zn.ptr(X.sum), zn.ptr(X.sumsq))
                # shortcut: average pool -> 1x1 float conv -> BatchNorm (with scale)
                self._call("zkb_avgpool2_f32_fwd", zn.ptr(self.h[u]), B, hi, wi, ci, zn.ptr(X.pool))
                self._call("zkb_dense_f32_fwd", zn.ptr(X.pool), zn.ptr(X.sc_kernel), 0, zn.ptr(X.t), rows, ci, co, co)
                self._call("zkb_bn_stats", zn.ptr(X.t), 0, rows, co, zn.ptr(X.sc_sum), zn.ptr(X.sc_sumsq))
                self._call("zkb_bn_gain_finalize", zn.ptr(X.sc_sum), zn.ptr(X.sc_sumsq), float(rows), zn.ptr(self.ones_c),
                           zn.ptr(X.sc_gamma), BN_EPS, self.momentum, 1, co, zn.ptr(X.sc_mean), zn.ptr(X.sc_rs), zn.ptr(X.sc_gain),
                           mm(X.sc_mm), mm(X.sc_mv))
                self._call("zkb_bn_gain_add_f32", zn.ptr(X.t), rows, co, zn.ptr(X.sc_mean), zn.ptr(X.sc_gain), zn.ptr(X.sc_beta),
                           0, zn.ptr(X.sc))
                shortcut = X.sc
            self._call("zkb_bn_gain_finalize", zn.ptr(X.sum), zn.ptr(X.sumsq), float(rows), zn.ptr(X.alpha), zn.ptr(X.gamma),
                       BN_EPS, self.momentum, 1, co, zn.ptr(X.mean_y), zn.ptr(X.rs_bwd), zn.ptr(X.gain), mm(X.mm), mm(X.mv))
            self._call("zkb_bn_gain_add", zn.ptr(X.y), rows, co, zn.ptr(X.mean_y), zn.ptr(X.gain), zn.ptr(X.beta),
                       zn.ptr(shortcut), zn.ptr(self.h[u + 1]))
        oh, ow = self.out_hw
        self._call("zkb_global_avgpool_fwd", zn.ptr(self.h[self.U]), B, oh * ow, self.out_c, zn.ptr(self.pooled))
        self._call("zkb_dense_f32_fwd", zn.ptr(self.pooled), zn.ptr(self.dense_w), zn.ptr(self.dense_b), zn.ptr(self.logits),
                   B, self.out_c, self.NC, self.NC)

    def _bn_backward_sums(self, d, x, is_i16, rows, co, mean, rs, beta, sum_d, sum_dx, dbeta, dgamma, mean_d, mean_dx):
        """BatchNorm (with scale) backward, first half: sums with the unmasked gradient, dbeta, the two
        means, and dgamma = sum d*xhat (the same finalize with the sums swapped; its means are scratch)."""
        self._call("zkb_bn_bwd_reduce", zn.ptr(d), zn.ptr(x), is_i16, rows, co, zn.ptr(mean), zn.ptr(rs), zn.ptr(beta), 0,
                   zn.ptr(sum_d), zn.ptr(sum_dx))
        self._call("zkb_bn_bwd_finalize", zn.ptr(sum_d), zn.ptr(sum_dx), float(rows), 1.0, co, zn.ptr(dbeta), zn.ptr(mean_d), zn.ptr(mean_dx))
        self._call("zkb_bn_bwd_finalize", zn.ptr(sum_dx), zn.ptr(sum_d), float(rows), 1.0, co, zn.ptr(dgamma),
                   zn.ptr(self.scratch_c), self.scratch_c[co:].data_ptr())

    def backward(self):
        B, NC = self.B, self.NC
        oh, ow = self.out_hw
        self._call("zkb_softmax_ce", zn.ptr(self.logits), NC, zn.ptr(self.labels), B, NC, 1.0 / (B * self.world), zn.ptr(self.dlogits),
                   self.acc[self.loss_off:].data_ptr(), zn.ptr(self.correct))
        self._call("zkb_dense_f32_bwd", zn.ptr(self.pooled), zn.ptr(self.dense_w), zn.ptr(self.dlogits), NC, zn.ptr(self.dpooled),
                   zn.ptr(self.d_dense_w), zn.ptr(self.d_dense_b), B, self.out_c, NC)
        self._call("zkb_global_avgpool_bwd", zn.ptr(self.dpooled), B, oh * ow, self.out_c, zn.ptr(self.dh[self.U]))
        for u in range(self.U - 1, -1, -1):
            X = self.units[u]
            (hi, wi, ci), (ho, wo, co) = X.in_shape, X.out_shape
            rows_in, rows = B * hi * wi, B * ho * wo
            d = self.dh[u + 1]
            self._bn_backward_sums(d, X.y, 1, rows, co, X.mean_y, X.rs_bwd, X.beta, X.sum_d, X.sum_dx, X.dbeta, X.dgamma,
                                   X.mean_d, X.mean_dx)
            self._call("zkb_bn_gain_bwd_apply", zn.ptr(d), zn.ptr(X.y), rows, co, zn.ptr(X.mean_y), zn.ptr(X.rs_bwd), zn.ptr(X.gain),
                       zn.ptr(X.mean_d), zn.ptr(X.mean_dx), zn.ptr(X.dy_hi), zn.ptr(X.dy_lo))
            # dy = alpha*dz, so the wgrad output IS the latent-kernel gradient alpha * wgrad(sign(h), dz);
            # the |w| <= 1 mask of magnitude_aware_sign's ste is applied by the Adam kernel
            if X.kind == "plain":
                self._call("zkb_bconv2d_dgrad", zn.ptr(X.dy_hi), zn.ptr(X.dy_lo), zn.ptr(X.w_dgrad), zn.ptr(X.ds), B, hi, wi, ci, co, 3)
                self._call("zkb_bconv2d_wgrad", zn.ptr(X.s), zn.ptr(X.dy_hi), zn.ptr(X.dy_lo), zn.ptr(X.dkernel), B, hi, wi, ci, co, 3)
                d_short = d
            else:
                self._call("zkb_dilate2_hi_lo", zn.ptr(X.dy_hi), zn.ptr(X.dy_lo), B, hi, wi, co, 3, zn.ptr(X.dil_hi), zn.ptr(X.dil_lo))
                self._call("zkb_bconv2d_dgrad", zn.ptr(X.dil_hi), zn.ptr(X.dil_lo), zn.ptr(X.w_dgrad), zn.ptr(X.ds), B, hi, wi, ci, co, 3)
                self._call("zkb_bconv2d_wgrad_s2", zn.ptr(X.s), zn.ptr(X.dy_hi), zn.ptr(X.dy_lo), zn.ptr(X.dkernel), B, hi, wi, ci, co, 3)
                # shortcut: BatchNorm -> 1x1 conv -> average pool, backwards
                self._bn_backward_sums(d, X.t, 0, rows, co, X.sc_mean, X.sc_rs, X.sc_beta, X.sc_sum_d, X.sc_sum_dx,
                                       X.d_sc_beta, X.d_sc_gamma, X.sc_mean_d, X.sc_mean_dx)
                self._call("zkb_bn_gain_bwd_apply_f32", zn.ptr(d), zn.ptr(X.t), rows, co, zn.ptr(X.sc_mean), zn.ptr(X.sc_rs),
                           zn.ptr(X.sc_gain), zn.ptr(X.sc_mean_d), zn.ptr(X.sc_mean_dx), zn.ptr(X.dt))
                self._call("zkb_dense_f32_bwd", zn.ptr(X.pool), zn.ptr(X.sc_kernel), zn.ptr(X.dt), co, zn.ptr(X.dpool),
                           zn.ptr(X.d_sc_kernel), 0, rows, ci, co)
                self._call("zkb_avgpool2_f32_bwd", zn.ptr(X.dpool), B, hi, wi, ci, zn.ptr(X.dh_sc))
                d_short = X.dh_sc
            # approx_sign backward, plus the gradient that arrives over the float shortcut
            self._call("zkb_approx_sign_bwd_add", zn.ptr(X.ds), zn.ptr(self.h[u]), zn.ptr(d_short), zn.ptr(self.dh[u]), rows_in * ci)

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
        self.dp.all_reduce_sum_(self.grads)
        if apply_update:
            self.optimizer_step()
        self.launches_per_step = self._n_launch

    def last_loss(self) -> float:
        """Mean loss over THIS replica's samples of the last step."""
        return float(self.acc[self.loss_off].item()) / self.B

    def last_accuracy(self) -> float:
        return float(self.correct.item()) / self.B


def bireal18_plan():
    """Unit kinds of larq-zoo's Bi-Real-Net-18 after the stem: four 64-channel units, then three
    stages (128, 256, 512 channels) of a down-sampling unit followed by three plain ones."""
    return ["plain"] * 4 + (["down"] + ["plain"] * 3) * 3


class BiRealNetEngine:
    """The whole zoo model: stem (float Conv2D(filters, 7, strides=2, "same", use_bias=False) ->
    BatchNormalization -> MaxPool2D(3, 2, "same")) in front of a ``BiRealUnitsEngine``.  Images are
    float32 [B,H,W,3] (what ``zkb_preprocess_u8`` writes).  The stem kernel is a plain float Conv2D:
    Adam without the latent-weight clip."""

    def __init__(self, units, image_hw, batch_size: int, num_classes: int, stem_filters: int = 64,
                 stem_ksize: int = 7, stem_stride: int = 2, **kw):
        zn.load()
        if not torch.cuda.is_available():
            raise RuntimeError("BiRealNetEngine needs a CUDA device (B200); there is no CPU fallback")
        if (stem_ksize, stem_stride) not in ((7, 2), (11, 4)):
            raise ValueError("stem: (ksize, stride) is (7, 2) or (11, 4)")
        if stem_filters % 64 != 0:
            raise ValueError(f"stem_filters={stem_filters} must be a multiple of 64")
        H, W = int(image_hw[0]), int(image_hw[1])
        self.H, self.W, self.B, self.F = H, W, int(batch_size), int(stem_filters)
        self.ks, self.stride = int(stem_ksize), int(stem_stride)
        self.H1, self.W1 = -(-H // self.stride), -(-W // self.stride)      # conv output ('same')
        self.H2, self.W2 = (self.H1 + 1) // 2, (self.W1 + 1) // 2           # pooled
        self.body = BiRealUnitsEngine(units, self.F, (self.H2, self.W2), batch_size, num_classes, **kw)
        dev, f32, B, F = self.body.device, torch.float32, self.B, self.F
        self.device = dev
        nk = self.ks * self.ks * 3 * F
        n = nk + 2 * F
        self.params = torch.zeros(n, dtype=f32, device=dev)
        self.grads = torch.zeros(n, dtype=f32, device=dev)
        self.adam_m = torch.zeros(n, dtype=f32, device=dev)
        self.adam_v = torch.zeros(n, dtype=f32, device=dev)
        self.kernel, self.dkernel = self.params[:nk], self.grads[:nk]
        self.gamma, self.dgamma = self.params[nk:nk + F], self.grads[nk:nk + F]
        self.beta, self.dbeta = self.params[nk + F:], self.grads[nk + F:]
        self.gamma.fill_(1.0)
        self.acc = torch.zeros(4 * F, dtype=torch.float64, device=dev)
        self.sum, self.sumsq, self.sum_d, self.sum_dx = (self.acc[i * F:(i + 1) * F] for i in range(4))
        self.mean, self.rs, self.gain, self.mean_d, self.mean_dx = (torch.zeros(F, dtype=f32, device=dev) for _ in range(5))
        self.mm, self.mv = torch.zeros(F, dtype=f32, device=dev), torch.ones(F, dtype=f32, device=dev)
        self.ones = torch.ones(F, dtype=f32, device=dev)
        self.scratch = torch.zeros(2 * F, dtype=f32, device=dev)
        self.x = torch.zeros(B, H, W, 3, dtype=f32, device=dev)
        self.z = torch.zeros(B, self.H1, self.W1, F, dtype=f32, device=dev)     # conv output
        self.a = torch.zeros(B, self.H1, self.W1, F, dtype=f32, device=dev)     # BatchNorm(z)
        self.da = torch.zeros(B, self.H1, self.W1, F, dtype=f32, device=dev)
        self.dz = torch.zeros(B, self.H1, self.W1, F, dtype=f32, device=dev)
        self.arg = torch.zeros(B, self.H2, self.W2, F, dtype=torch.uint8, device=dev)

    # ------------------------------------------------------------------ parameters in / out
    def load_params(self, p: Dict[str, np.ndarray]):
        """The body's names (``BiRealUnitsEngine.load_params``) plus ``stem_kernel`` HWIO [k,k,3,F],
        ``stem_gamma``, ``stem_beta`` [F]."""
        self.body.load_params(p)
        self.kernel.copy_(torch.as_tensor(np.ascontiguousarray(p["stem_kernel"]), dtype=torch.float32).reshape(-1))
        self.gamma.copy_(torch.as_tensor(p["stem_gamma"], dtype=torch.float32))
        self.beta.copy_(torch.as_tensor(p["stem_beta"], dtype=torch.float32))

    def _stem_named(self, t):
        nk, F = self.ks * self.ks * 3 * self.F, self.F
        t = t.detach().cpu().numpy()
        return {"stem_kernel": t[:nk].reshape(self.ks, self.ks, 3, F).copy(), "stem_gamma": t[nk:nk + F].copy(),
                "stem_beta": t[nk + F:].copy()}

    def get_params(self) -> Dict[str, np.ndarray]:
        return {**self.body.get_params(), **self._stem_named(self.params)}

    def get_grads(self) -> Dict[str, np.ndarray]:
        out = {**self.body.get_grads(), **self._stem_named(self.grads)}
        out.pop("h0")           # the gradient entering the stem's pool; not a parameter
        return out

    # ------------------------------------------------------------------ one step
    def stem_forward(self, training: bool = True):
        c, B, F = self.body._call, self.B, self.F
        rows = B * self.H1 * self.W1
        mm = (lambda t: zn.ptr(t)) if training else (lambda t: 0)
        c("zkb_conv_stem_f32_fwd", zn.ptr(self.x), zn.ptr(self.kernel), zn.ptr(self.z), B, self.H, self.W, 3, F, self.ks, self.stride)
        c("zkb_bn_stats", zn.ptr(self.z), 0, rows, F, zn.ptr(self.sum), zn.ptr(self.sumsq))
        c("zkb_bn_gain_finalize", zn.ptr(self.sum), zn.ptr(self.sumsq), float(rows), zn.ptr(self.ones), zn.ptr(self.gamma),
          BN_EPS, self.body.momentum, 1, F, zn.ptr(self.mean), zn.ptr(self.rs), zn.ptr(self.gain), mm(self.mm), mm(self.mv))
        c("zkb_bn_gain_add_f32", zn.ptr(self.z), rows, F, zn.ptr(self.mean), zn.ptr(self.gain), zn.ptr(self.beta), 0, zn.ptr(self.a))
        c("zkb_maxpool3s2_f32_fwd", zn.ptr(self.a), B, self.H1, self.W1, F, zn.ptr(self.body.h[0]), zn.ptr(self.arg))

    def stem_backward(self):
        c, B, F = self.body._call, self.B, self.F
        rows = B * self.H1 * self.W1
        c("zkb_maxpool3s2_f32_bwd", zn.ptr(self.body.dh[0]), zn.ptr(self.arg), B, self.H1, self.W1, F, zn.ptr(self.da))
        c("zkb_bn_bwd_reduce", zn.ptr(self.da), zn.ptr(self.z), 0, rows, F, zn.ptr(self.mean), zn.ptr(self.rs), zn.ptr(self.beta), 0,
          zn.ptr(self.sum_d), zn.ptr(self.sum_dx))
        c("zkb_bn_bwd_finalize", zn.ptr(self.sum_d), zn.ptr(self.sum_dx), float(rows), 1.0, F, zn.ptr(self.dbeta), zn.ptr(self.mean_d),
          zn.ptr(self.mean_dx))
        c("zkb_bn_bwd_finalize", zn.ptr(self.sum_dx), zn.ptr(self.sum_d), float(rows), 1.0, F, zn.ptr(self.dgamma), zn.ptr(self.scratch),
          self.scratch[F:].data_ptr())
        c("zkb_bn_gain_bwd_apply_f32", zn.ptr(self.da), zn.ptr(self.z), rows, F, zn.ptr(self.mean), zn.ptr(self.rs), zn.ptr(self.gain),
          zn.ptr(self.mean_d), zn.ptr(self.mean_dx), zn.ptr(self.dz))
        c("zkb_conv_stem_f32_wgrad", zn.ptr(self.x), zn.ptr(self.dz), zn.ptr(self.dkernel), B, self.H, self.W, 3, F, self.ks, self.stride)

    def train_step(self, images: torch.Tensor, labels: torch.Tensor, apply_update: bool = True):
        """images: float32 [B,H,W,3]; labels: integer [B]."""
        body = self.body
        body._stream = torch.cuda.current_stream(self.device).cuda_stream
        body._n_launch = 0
        for t in (self.acc, self.grads, body.acc, body.correct, body.grads):
            t.zero_()
        if images.data_ptr() != self.x.data_ptr():
            self.x.copy_(images)
        body.labels.copy_(labels.to(torch.int32))
        body.prepare_weights()
        self.stem_forward(training=True)
        body.forward(training=True)
        body.backward()
        self.stem_backward()
        body.dp.all_reduce_sum_(body.grads)
        body.dp.all_reduce_sum_(self.grads)
        if apply_update:
            body.optimizer_step()
            body._call("zkb_adam_clip", zn.ptr(self.params), zn.ptr(self.grads), zn.ptr(self.adam_m), zn.ptr(self.adam_v),
                       self.params.numel(), body._lr_t(body.step_count), body.b1, body.b2, body.eps, 1.0, 0, 0)
        self.launches_per_step = body._n_launch

    def last_loss(self) -> float:
        return self.body.last_loss()

    def last_accuracy(self) -> float:
        return self.body.last_accuracy()
