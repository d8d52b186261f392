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

What is NOT here: the float 7x7/2 stem conv and a Keras functional-API front end -- the stack takes
float activations [B,H,W,C] as input (``zkb_maxpool3s2_f32_*``, the stem's pool, is available as a
kernel).  The 1x1 float conv of the down-sampling shortcut runs on plain SIMT kernels.
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
    """A stack of binary residual units on a [B, H, W, channels] float map, then global average
    pooling and a float classifier.  ``units``: a count of plain units, or a sequence of "plain" /
    "down".  Channel counts must be multiples of 128 (the int8 forward GEMM reads 128 channels per
    shared-memory row); "down" needs an even map."""

    def __init__(self, units, channels: int, hw, batch_size: int, num_classes: int,
                 learning_rate: float = 5e-3, bn_momentum: float = 0.99, beta_1: float = 0.9,
                 beta_2: float = 0.999, epsilon: float = 1e-7, device: Optional[torch.device] = None):
        zn.load()
        if not torch.cuda.is_available():
            raise RuntimeError("BiRealUnitsEngine needs a CUDA device (B200); there is no CPU fallback")
        self.plan = ["plain"] * units if isinstance(units, int) else [str(u) for u in units]
        if not self.plan or any(u not in ("plain", "down") for u in self.plan):
            raise ValueError('units: a positive count or a sequence of "plain" / "down"')
        if channels % 128 != 0:
            raise ValueError(f"channels={channels} must be a multiple of 128")
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
            if X.kind == "plain":
                self._call("zkb_bconv2d_fwd_i8_stats", zn.ptr(X.s8), zn.ptr(X.w_fwd8), zn.ptr(X.y), B, hi, wi, ci, co, 3,
                           zn.ptr(X.sum), zn.ptr(X.sumsq))
                shortcut = self.h[u]
            else:
                self._call("zkb_bconv2d_fwd_i8_s2", zn.ptr(X.s8), zn.ptr(X.w_fwd8), zn.ptr(X.y), B, hi, wi, ci, co, 3)
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
        self._call("zkb_softmax_ce", zn.ptr(self.logits), NC, zn.ptr(self.labels), B, NC, 1.0 / B, zn.ptr(self.dlogits),
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
        if apply_update:
            self.optimizer_step()
        self.launches_per_step = self._n_launch

    def last_loss(self) -> float:
        return float(self.acc[self.loss_off].item()) / self.B

    def last_accuracy(self) -> float:
        return float(self.correct.item()) / self.B
