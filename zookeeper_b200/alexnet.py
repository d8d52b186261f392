"""BinaryAlexNet train step on libzkb200 (BASELINE.json configs[3]).

larq-zoo's BinaryAlexNet is BinaryNet's recipe at ImageNet shape -- ``QuantConv2D`` / ``QuantDense``
with ``ste_sign`` input and kernel quantizers and ``weight_clip``, ``BatchNormalization(scale=False)``
behind every layer, Adam on the latent weights -- with other layer shapes:

    conv 11x11/4 (64, float image x sign(W)) -> pool -> BN     conv 5x5 (192) -> pool -> BN
    conv 3x3 (384) -> BN    conv 3x3 (384) -> BN    conv 3x3 (256) -> pool -> BN
    Flatten -> dense 4096 -> BN -> dense 4096 -> BN -> dense classes -> BN -> softmax

(pool = MaxPool2D(3, 2), "valid", BEFORE the BatchNormalization: it pools the integer conv output).
The reference reaches such a model through the hook its example uses for BinaryNet, a ``@factory``
model built from larq layers (/root/reference/examples/larq_experiment.py:41-57, 59-103; compiled and
fitted at :117-122, :141-153); the model definition itself is larq-zoo's, not in the reference tree,
restated in oracle/alexnet_oracle.py.

Everything arithmetic is a C-ABI call (``include/zkb200.h``).  The binary layers run on the tcgen05
GEMMs of the BinaryNet path: int8 forward where the input width is a multiple of 128, the bf16 +-1
forward for the 64- and 192-channel inputs (same integers), bf16 hi/lo backward; the weight-gradient
GEMM tiles 64-channel units (``csrc/gemm_tc.cu``, ``WGemmParams::flat64``), output widths that are
not a multiple of the MMA tile (192, 384) are clipped by TMA.  New kernels: the float stem
(``csrc/stem.cu``) and the overlapping pool (``csrc/pool3.cu``); the per-channel BatchNorm kernels walk
the 4096-wide dense maps in column slices.  No CPU or PyTorch fallback: PyTorch holds device memory.
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import numpy as np
import torch

from . import native as zn
from .dataparallel import DataParallel

BN_EPS = 1e-3
FEATURES = (64, 192, 384, 384, 256)
KSIZES = (11, 5, 3, 3, 3)
STRIDES = (4, 1, 1, 1, 1)
POOLS = (True, True, False, False, True)


class _Layer:
    pass


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class BinaryAlexNetEngine:
    """images: float32 [B,H,W,3] in [-1, 1] (what ``zkb_preprocess_u8`` writes); labels: integer [B]."""

    def __init__(self, image_hw, batch_size: int, num_classes: int, features: Sequence[int] = FEATURES,
                 dense_units: int = 4096, learning_rate: float = 5e-3, bn_momentum: float = 0.9, beta_1: float = 0.9,
                 beta_2: float = 0.999, epsilon: float = 1e-7, device: Optional[torch.device] = None,
                 data_parallel: Optional[DataParallel] = None):
        zn.load()
        if not torch.cuda.is_available():
            raise RuntimeError("BinaryAlexNetEngine needs a CUDA device (B200); there is no CPU fallback")
        # one process per GPU: the batch shards by sample, BatchNorm statistics stay per replica, the
        # flat gradient arena is summed over the replicas (the dense kernels -- 95 % of it -- while
        # the conv layers are still in their backward pass)
        self.dp = data_parallel if data_parallel is not None else DataParallel()
        self.world = self.dp.world
        self._side = None
        if len(features) != 5 or any(f % 64 for f in features):
            raise ValueError("features: five channel counts, multiples of 64")
        if dense_units % 128:
            raise ValueError("dense_units must be a multiple of 128")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.H, self.W, self.B, self.NC = int(image_hw[0]), int(image_hw[1]), int(batch_size), int(num_classes)
        self.features, self.dense_units = tuple(int(f) for f in features), int(dense_units)
        self.lr, self.b1, self.b2, self.eps, self.momentum = float(learning_rate), beta_1, beta_2, epsilon, float(bn_momentum)
        self.step_count = 0
        self._allocate()

    # ------------------------------------------------------------------ memory
    def _allocate(self):
        dev, B = self.device, self.B
        f32, i16, bf16, u8, i8 = torch.float32, torch.int16, torch.bfloat16, torch.uint8, torch.int8
        specs = []                       # (kind, cin, cout, ksize, stride, pool, in_h, in_w, conv_h, conv_w, out_h, out_w)
        h, w, c = self.H, self.W, 3
        for f, k, s, pool in zip(self.features, KSIZES, STRIDES, POOLS):
            ch, cw = -(-h // s), -(-w // s)
            oh, ow = ((ch - 3) // 2 + 1, (cw - 3) // 2 + 1) if pool else (ch, cw)
            if ch < 3 or cw < 3 or oh < 1 or ow < 1:
                raise ValueError(f"image {self.H}x{self.W} is too small for the three 3x3/2 pools")
            specs.append(("conv", c, f, k, s, pool, h, w, ch, cw, oh, ow))
            h, w, c = oh, ow, f
        flat = h * w * c
        if flat % 128:
            raise ValueError(f"flattened width {flat} must be a multiple of 128")
        for kin, kout in ((flat, self.dense_units), (self.dense_units, self.dense_units), (self.dense_units, self.NC)):
            specs.append(("dense", kin, kout, 1, 1, False, 1, 1, 1, 1, 1, 1))
        self.flat = flat
        cps = [s[2] if s[2] % 64 == 0 else _round_up(s[2], 64) for s in specs]
        k_sizes = [s[3] * s[3] * s[1] * s[2] for s in specs]
        # arena order [dense kernels | conv kernels | betas]: one clipped Adam launch over the kernels, one
        # plain over the betas; the gradients that are complete FIRST in the backward pass (the dense
        # kernels) are a prefix, so the early bucket of the gradient all-reduce is a plain slice
        order = [5, 6, 7, 0, 1, 2, 3, 4]
        k_off = np.zeros(len(specs), dtype=np.int64)
        pos = 0
        for i in order:
            k_off[i] = pos
            pos += _round_up(k_sizes[i], 4)
        self.n_dense_kernel = int(k_off[0])
        b_off = np.concatenate([[0], np.cumsum(cps)]).astype(np.int64)
        self.n_kernel, self.n_beta = int(pos), int(b_off[-1])
        n = self.n_kernel + self.n_beta
        self.params = torch.zeros(n, dtype=f32, device=dev)
        self.grads = torch.zeros(n, dtype=f32, device=dev)
        self.adam_m = torch.zeros(n, dtype=f32, device=dev)
        self.adam_v = torch.zeros(n, dtype=f32, device=dev)
        self.acc = torch.zeros(4 * self.n_beta + 2, dtype=torch.float64, device=dev)
        self.loss_off = 4 * self.n_beta
        self.stat = torch.zeros(4 * self.n_beta, dtype=f32, device=dev)
        self.moving_mean = torch.zeros(self.n_beta, dtype=f32, device=dev)
        self.moving_var = torch.ones(self.n_beta, dtype=f32, device=dev)
        self.correct = torch.zeros(1, dtype=torch.int32, device=dev)
        self.labels = torch.zeros(B, dtype=torch.int32, device=dev)
        self.x = torch.zeros(B, self.H, self.W, 3, dtype=f32, device=dev)
        self.L = []
        nb = self.n_beta
        for i, (kind, cin, cout, k, s, pool, ih, iw, ch, cw, oh, ow) in enumerate(specs):
            L = _Layer()
            L.idx, L.kind, L.cin, L.cout, L.k, L.stride, L.pool = i, kind, cin, cout, k, s, pool
            L.in_hw, L.conv_hw, L.out_hw, L.cp = (ih, iw), (ch, cw), (oh, ow), cps[i]
            L.last = i == len(specs) - 1
            ko, bo, cp = int(k_off[i]), self.n_kernel + int(b_off[i]), cps[i]
            L.kernel, L.dkernel = self.params[ko:ko + k_sizes[i]], self.grads[ko:ko + k_sizes[i]]
            L.beta, L.dbeta = self.params[bo:bo + cp], self.grads[bo:bo + cp]
            so = int(b_off[i])
            a = self.acc[4 * so:4 * so + 4 * cp]
            L.sum, L.sumsq, L.sum_da, L.sum_dax = a[:cp], a[cp:2 * cp], a[2 * cp:3 * cp], a[3 * cp:]
            L.mean, L.rstd = self.stat[so:so + cp], self.stat[nb + so:nb + so + cp]
            L.mean_da, L.mean_dax = self.stat[2 * nb + so:2 * nb + so + cp], self.stat[3 * nb + so:3 * nb + so + cp]
            L.mm, L.mv = self.moving_mean[so:so + cp], self.moving_var[so:so + cp]
            if kind == "conv":
                L.rows = B * oh * ow
                ydt = f32 if i == 0 else i16
                L.y = torch.zeros(B, ch, cw, cout, dtype=ydt, device=dev)          # conv output (layer 0: float)
                if pool:
                    L.p = torch.zeros(B, oh, ow, cout, dtype=ydt, device=dev)
                    L.arg = torch.zeros(B, oh, ow, cout, dtype=u8, device=dev)
                    L.dp = torch.zeros(B, oh, ow, cout, dtype=f32, device=dev)     # gradient w.r.t. the pooled map
                L.s = torch.zeros(B, oh, ow, cout, dtype=bf16, device=dev)
                L.s8 = torch.zeros(B, oh, ow, cout, dtype=i8, device=dev)
                L.ds = torch.zeros(B, oh, ow, cout, dtype=f32, device=dev)
                if i == 0:
                    L.wq = torch.zeros(k * k * cin * cout, dtype=f32, device=dev)
                    L.dz = torch.zeros(B, ch, cw, cout, dtype=f32, device=dev)
                else:
                    L.w_fwd = torch.zeros(cout, k * k * cin, dtype=bf16, device=dev)
                    L.w_fwd8 = torch.zeros(cout, k * k * cin, dtype=i8, device=dev)
                    L.w_dgrad = torch.zeros(cin, k * k * cout, dtype=bf16, device=dev)
                    L.dy_hi = torch.zeros(B, ch, cw, cout, dtype=bf16, device=dev)
                    L.dy_lo = torch.zeros(B, ch, cw, cout, dtype=bf16, device=dev)
            else:
                L.rows = B
                L.y = torch.zeros(B, cp, dtype=i16, device=dev)
                L.ds = torch.zeros(B, cp, dtype=f32, device=dev)
                if L.last:
                    L.logits = torch.zeros(B, cp, dtype=f32, device=dev)
                else:
                    L.s = torch.zeros(B, cp, dtype=bf16, device=dev)
                    L.s8 = torch.zeros(B, cp, dtype=i8, device=dev)
                L.w_fwd = torch.zeros(cp, cin, dtype=bf16, device=dev)
                L.w_fwd8 = torch.zeros(cp, cin, dtype=i8, device=dev)
                L.w_dgrad = torch.zeros(cin, cp, dtype=bf16, device=dev)
                L.dy_hi = torch.zeros(B, cp, dtype=bf16, device=dev)
                L.dy_lo = torch.zeros(B, cp, dtype=bf16, device=dev)
            self.L.append(L)

    # ------------------------------------------------------------------ parameters in / out
    def _kshape(self, L):
        return (L.k, L.k, L.cin, L.cout) if L.kind == "conv" else (L.cin, L.cout)

    def load_params(self, p: Dict[str, np.ndarray]):
        """``kernel{i}``: HWIO [k,k,Cin,Cout] (conv) or [K,N] (dense); ``beta{i}`` [Cout]; i = 0..7."""
        for L in self.L:
            L.kernel.copy_(torch.as_tensor(np.ascontiguousarray(p[f"kernel{L.idx}"]), dtype=torch.float32).reshape(-1))
            L.beta.zero_()
            L.beta[:L.cout].copy_(torch.as_tensor(p[f"beta{L.idx}"], dtype=torch.float32))

    def _named(self, kattr, battr):
        out = {}
        for L in self.L:
            out[f"kernel{L.idx}"] = getattr(L, kattr).detach().cpu().numpy().reshape(self._kshape(L)).copy()
            out[f"beta{L.idx}"] = getattr(L, battr)[:L.cout].detach().cpu().numpy().copy()
        return out

    def get_params(self) -> Dict[str, np.ndarray]:
        return self._named("kernel", "beta")

    def get_grads(self) -> Dict[str, np.ndarray]:
        return self._named("dkernel", "dbeta")

    # ------------------------------------------------------------------ one step
    def _call(self, name, *args):
        self._n_launch += 1
        zn.call(name, *args, self._stream)

    def prepare_weights(self):
        for L in self.L:
            if L.idx == 0:
                self._call("zkb_sign_f32", zn.ptr(L.kernel), zn.ptr(L.wq), L.kernel.numel())
            elif L.kind == "conv":
                self._call("zkb_prep_conv_weights", zn.ptr(L.kernel), L.k, L.cin, L.cout, zn.ptr(L.w_fwd), zn.ptr(L.w_dgrad), zn.ptr(L.w_fwd8))
            else:
                self._call("zkb_prep_dense_weights", zn.ptr(L.kernel), L.cin, L.cout, L.cp, zn.ptr(L.w_fwd), zn.ptr(L.w_dgrad), zn.ptr(L.w_fwd8))

    def forward(self, training: bool = True):
        B = self.B
        for L in self.L:
            prev = self.L[L.idx - 1] if L.idx else None
            stats_done = False
            if L.kind == "conv":
                (ih, iw), (ch, cw) = L.in_hw, L.conv_hw
                if L.idx == 0:
                    self._call("zkb_conv_stem_f32_fwd", zn.ptr(self.x), zn.ptr(L.wq), zn.ptr(L.y), B, ih, iw, 3, L.cout, L.k, L.stride)
                elif L.cin % 128 == 0:
                    if L.pool:
                        self._call("zkb_bconv2d_fwd_i8", zn.ptr(prev.s8), zn.ptr(L.w_fwd8), zn.ptr(L.y), B, ih, iw, L.cin, L.cout, L.k)
                    else:
                        self._call("zkb_bconv2d_fwd_i8_stats", zn.ptr(prev.s8), zn.ptr(L.w_fwd8), zn.ptr(L.y), B, ih, iw, L.cin, L.cout,
                                   L.k, zn.ptr(L.sum), zn.ptr(L.sumsq))
                        stats_done = True
                else:       # 64 / 192 input channels: the int8 GEMM reads 128 channels per shared-memory row
                    self._call("zkb_bconv2d_fwd", zn.ptr(prev.s), zn.ptr(L.w_fwd), zn.ptr(L.y), B, ih, iw, L.cin, L.cout, L.k)
                is_i16 = 0 if L.idx == 0 else 1
                xbn = L.y
                if L.pool:
                    self._call("zkb_maxpool3s2_fwd", zn.ptr(L.y), is_i16, B, ch, cw, L.cout, 0, zn.ptr(L.p), zn.ptr(L.arg))
                    xbn = L.p
                bessel = 1
            else:
                self._call("zkb_bdense_fwd_i8_stats", zn.ptr(prev.s8), zn.ptr(L.w_fwd8), zn.ptr(L.y), L.cp, B, L.cin, L.cp,
                           zn.ptr(L.sum), zn.ptr(L.sumsq))
                stats_done, is_i16, xbn, bessel = True, 1, L.y, 0
            if not stats_done:
                self._call("zkb_bn_stats", zn.ptr(xbn), is_i16, L.rows, L.cp, zn.ptr(L.sum), zn.ptr(L.sumsq))
            mm = (zn.ptr(L.mm), zn.ptr(L.mv)) if training else (0, 0)
            self._call("zkb_bn_finalize", zn.ptr(L.sum), zn.ptr(L.sumsq), float(L.rows), BN_EPS, self.momentum, bessel, L.cp,
                       zn.ptr(L.mean), zn.ptr(L.rstd), *mm)
            self._call("zkb_bn_apply_sign", zn.ptr(xbn), is_i16, L.rows, L.cp, zn.ptr(L.mean), zn.ptr(L.rstd), zn.ptr(L.beta),
                       0 if L.last else zn.ptr(L.s), zn.ptr(L.logits) if L.last else 0, 0 if L.last else zn.ptr(L.s8))

    def backward(self):
        B = self.B
        last = self.L[-1]
        self._call("zkb_softmax_ce", zn.ptr(last.logits), last.cp, zn.ptr(self.labels), B, self.NC, 1.0 / (B * self.world),
                   zn.ptr(last.ds), self.acc[self.loss_off:].data_ptr(), zn.ptr(self.correct))
        self._pending = []
        for L in reversed(self.L):
            prev = self.L[L.idx - 1] if L.idx else None
            is_i16 = 0 if L.idx == 0 else 1
            xbn = L.p if (L.kind == "conv" and L.pool) else L.y
            ste = 0 if L.last else 1
            self._call("zkb_bn_bwd_reduce", zn.ptr(L.ds), zn.ptr(xbn), is_i16, L.rows, L.cp, zn.ptr(L.mean), zn.ptr(L.rstd),
                       zn.ptr(L.beta), ste, zn.ptr(L.sum_da), zn.ptr(L.sum_dax))
            self._call("zkb_bn_bwd_finalize", zn.ptr(L.sum_da), zn.ptr(L.sum_dax), float(L.rows), 1.0, L.cp, zn.ptr(L.dbeta),
                       zn.ptr(L.mean_da), zn.ptr(L.mean_dax))
            common = (zn.ptr(L.ds), zn.ptr(xbn), is_i16, L.rows, L.cp, zn.ptr(L.mean), zn.ptr(L.rstd), zn.ptr(L.beta),
                      zn.ptr(L.mean_da), zn.ptr(L.mean_dax), ste)
            if L.kind == "conv" and L.pool:
                ch, cw = L.conv_hw
                self._call("zkb_bn_bwd_apply", *common, 0, zn.ptr(L.dp), 0, 0, 0, 0, 0)
                if L.idx == 0:
                    self._call("zkb_maxpool3s2_bwd", zn.ptr(L.dp), zn.ptr(L.arg), B, ch, cw, L.cout, 0, zn.ptr(L.dz), 0, 0)
                else:
                    self._call("zkb_maxpool3s2_bwd", zn.ptr(L.dp), zn.ptr(L.arg), B, ch, cw, L.cout, 0, 0, zn.ptr(L.dy_hi), zn.ptr(L.dy_lo))
            else:
                self._call("zkb_bn_bwd_apply", *common, 1, 0, zn.ptr(L.dy_hi), zn.ptr(L.dy_lo), 0, 0, 0)
            if L.idx == 0:
                ih, iw = L.in_hw
                # gradient w.r.t. sign(W0); the |w| <= 1 mask of ste_sign is the Adam kernel's
                self._call("zkb_conv_stem_f32_wgrad", zn.ptr(self.x), zn.ptr(L.dz), zn.ptr(L.dkernel), B, ih, iw, 3, L.cout, L.k, L.stride)
            elif L.kind == "conv":
                ih, iw = L.in_hw
                self._call("zkb_bconv2d_dgrad", zn.ptr(L.dy_hi), zn.ptr(L.dy_lo), zn.ptr(L.w_dgrad), zn.ptr(prev.ds), B, ih, iw, L.cin, L.cout, L.k)
                self._call("zkb_bconv2d_wgrad", zn.ptr(prev.s), zn.ptr(L.dy_hi), zn.ptr(L.dy_lo), zn.ptr(L.dkernel), B, ih, iw, L.cin, L.cout, L.k)
            else:
                self._call("zkb_bdense_dgrad", zn.ptr(L.dy_hi), zn.ptr(L.dy_lo), zn.ptr(L.w_dgrad), zn.ptr(prev.ds), B, L.cin, L.cp)
                self._call("zkb_bdense_wgrad", zn.ptr(prev.s), zn.ptr(L.dy_hi), zn.ptr(L.dy_lo), zn.ptr(L.dkernel), L.cout, B, L.cin, L.cp, L.cout)
                if L.idx == 5 and self.world > 1:
                    # every dense kernel gradient has been enqueued: sum them over the replicas on a side
                    # stream, next to the conv layers' backward
                    main = torch.cuda.current_stream(self.device)
                    if self._side is None:
                        self._side = torch.cuda.Stream(self.device)
                        self._ev = torch.cuda.Event()
                    self._ev.record(main)
                    self._side.wait_event(self._ev)
                    with torch.cuda.stream(self._side):
                        self._pending.append(self.dp.all_reduce_sum_async_(self.grads[:self.n_dense_kernel]))

    def finish_gradient_reduce(self):
        """Sum the rest of the gradient arena over the replicas and join the early bucket."""
        if self.world == 1:
            return
        if self._pending:
            self._pending.append(self.dp.all_reduce_sum_async_(self.grads[self.n_dense_kernel:]))
        else:
            self._pending.append(self.dp.all_reduce_sum_async_(self.grads))
        for work in self._pending:
            if work is not None:
                work.wait()          # orders the current stream behind the collective; no host block
        self._pending = []

    def _lr_t(self, t: int) -> float:
        f = np.float32
        return float(f(self.lr) * np.sqrt(f(1) - f(self.b2) ** f(t), dtype=f) / (f(1) - f(self.b1) ** f(t)))

    def optimizer_step(self):
        self.step_count += 1
        lr_t, nk = self._lr_t(self.step_count), self.n_kernel
        self._call("zkb_adam_clip", zn.ptr(self.params), zn.ptr(self.grads), zn.ptr(self.adam_m), zn.ptr(self.adam_v), nk, lr_t,
                   self.b1, self.b2, self.eps, 1.0, 1, 0)
        self._call("zkb_adam_clip", self.params[nk:].data_ptr(), self.grads[nk:].data_ptr(), self.adam_m[nk:].data_ptr(),
                   self.adam_v[nk:].data_ptr(), self.n_beta, lr_t, self.b1, self.b2, self.eps, 1.0, 0, 0)

    def train_step(self, images: torch.Tensor, labels: torch.Tensor, apply_update: bool = True):
        self._stream = torch.cuda.current_stream(self.device).cuda_stream
        self._n_launch = 0
        self.acc.zero_()
        self.correct.zero_()
        self.grads.zero_()
        if images.data_ptr() != self.x.data_ptr():
            self.x.copy_(images)
        self.labels.copy_(labels.to(torch.int32))
        self.prepare_weights()
        self.forward(training=True)
        self.backward()
        self.finish_gradient_reduce()
        if apply_update:
            self.optimizer_step()
        self.launches_per_step = self._n_launch

    def last_loss(self) -> float:
        """Mean loss over THIS replica's samples of the last step."""
        return float(self.acc[self.loss_off].item()) / self.B

    def last_accuracy(self) -> float:
        return float(self.correct.item()) / self.B

    def logits(self) -> np.ndarray:
        return self.L[-1].logits[:, :self.NC].detach().cpu().numpy()
