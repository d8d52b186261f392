"""BinaryNet train-step engine: Python host code driving libzkb200.so through ctypes.

This is what replaces ``tf.keras.Sequential.fit`` for the model built by the reference's
``BinaryNet.build`` (examples/larq_experiment.py:51-103) compiled with
``sparse_categorical_crossentropy`` + ``Adam`` (:117-122, :141-153).  PyTorch is used for device
memory, streams and ``torch.distributed`` only; every arithmetic step is a hand-written sm_100a
kernel behind the C ABI (no CPU / PyTorch fallback: a missing library raises).

Data layout (all NHWC, channels innermost, resident in HBM for the whole run):
  x0        f32  [B,H,W,Cin]            preprocessed images
  y_i       f32 (first layer) / int16   pre-BatchNorm conv / dense output (exact integers)
  p_i, arg  int16 / u8                  2x2 max-pooled map and winning window slot
  s_i       bf16 (+-1)                  sign(BatchNorm(.)) -- operand of the next binary GEMM
  ds_i      f32                         gradient w.r.t. s_i (output of the next layer's dgrad)
  dy_i      bf16 hi + bf16 lo           gradient w.r.t. y_i, split so that tensor-core products
                                        with the exact +-1 operand keep ~16 mantissa bits
Parameters, gradients and Adam moments live in flat f32 arenas: [all kernels | all betas(padded)].
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import native as zn
from .dataparallel import DataParallel

BN_EPS = 1e-3
BN_MOMENTUM = 0.99
NPAD = 64  # dense output widths below 64 are padded to 64 columns inside the engine


@dataclass
class LayerSpec:
    """One QuantConv2D / QuantDense + (MaxPool2D) + BatchNormalization block."""

    kind: str            # "conv" | "dense"
    cin: int
    cout: int
    ksize: int = 3
    padding: str = "same"
    quant_input: bool = True
    pool: bool = False


def binarynet_layers(input_shape=(28, 28, 1), filters=128, dense_units=1024, num_classes=10,
                     kernel_size=3) -> List[LayerSpec]:
    """Topology of ``BinaryNet.build`` (examples/larq_experiment.py:59-103)."""
    if isinstance(kernel_size, (tuple, list)):
        if kernel_size[0] != kernel_size[1]:
            raise ValueError("only square kernels are supported")
        kernel_size = kernel_size[0]
    h, w, c = input_shape
    f = filters
    layers = [
        LayerSpec("conv", c, f, kernel_size, "valid", quant_input=False),
        LayerSpec("conv", f, f, kernel_size, "same", pool=True),
        LayerSpec("conv", f, 2 * f, kernel_size, "same"),
        LayerSpec("conv", 2 * f, 2 * f, kernel_size, "same", pool=True),
        LayerSpec("conv", 2 * f, 4 * f, kernel_size, "same"),
        LayerSpec("conv", 4 * f, 4 * f, kernel_size, "same", pool=True),
    ]
    h, w = h - kernel_size + 1, w - kernel_size + 1
    for _ in range(3):
        h, w = h // 2, w // 2
    layers += [
        LayerSpec("dense", h * w * 4 * f, dense_units, 1, "valid"),
        LayerSpec("dense", dense_units, dense_units, 1, "valid"),
        LayerSpec("dense", dense_units, num_classes, 1, "valid"),
    ]
    return layers


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def gradient_buckets(k_off, kbase: int, min_bytes: int = 8 << 20):
    """Slices of the flat f32 gradient arena ``[all betas | kernel 0 | kernel 1 | ...]`` in the order
    the backward pass completes them.  ``k_off[i]`` is kernel i's offset inside the kernel part
    (``k_off[-1]`` its length), ``kbase`` the length of the beta prefix.  Walking the layers from the
    last to the first, a bucket is closed once it holds ``min_bytes`` of kernel gradients; the final
    bucket takes what is left plus every beta gradient.  Each entry is (lowest layer index in the
    bucket, start, stop): the bucket is complete when that layer's weight-gradient GEMM is (the
    final one: when the whole backward pass is)."""
    n_layers = len(k_off) - 1
    buckets, stop = [], kbase + int(k_off[-1])
    for i in range(n_layers - 1, 1, -1):
        start = kbase + int(k_off[i])
        if (stop - start) * 4 >= min_bytes:
            buckets.append((i, start, stop))
            stop = start
    buckets.append((0, 0, stop))
    return buckets


class _Layer:
    """Device buffers of one block."""

    pass


class BinaryNetEngine:
    """Holds parameters, optimizer state and activations of a BinaryNet on one GPU and runs
    train / eval steps.  With ``world_size > 1`` (one process per GPU) gradients are summed with
    an NCCL all-reduce over the flat gradient arena; BatchNorm statistics are per replica unless
    ``sync_bn`` is set, in which case the per-channel sums are all-reduced as well."""

    def __init__(self, layers: Sequence[LayerSpec], input_shape: Tuple[int, int, int], batch_size: int,
                 num_classes: int, learning_rate: float = 5e-3, device: Optional[torch.device] = None,
                 process_group=None, sync_bn: bool = False, beta_1: float = 0.9, beta_2: float = 0.999,
                 epsilon: float = 1e-7, forward_int8: bool = True, data_parallel: bool = True,
                 share_state_with: Optional["BinaryNetEngine"] = None):
        """``share_state_with``: another engine of the same model (any batch size) whose parameter,
        Adam-moment and moving-statistics arenas and step counter this engine ALIASES instead of
        allocating its own -- one set of weights seen through several batch sizes (training at one
        batch size, validation at another) without copies that could go stale."""
        zn.load()  # fail loudly now if the CUDA library is missing
        if not torch.cuda.is_available():
            raise RuntimeError("BinaryNetEngine needs a CUDA device (B200); there is no CPU fallback")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.layers = list(layers)
        self.input_shape = tuple(input_shape)
        self.B = int(batch_size)
        self.num_classes = int(num_classes)
        self.lr, self.b1, self.b2, self.eps = float(learning_rate), beta_1, beta_2, epsilon
        self.pg = process_group
        self.dp = DataParallel(process_group, enabled=data_parallel)
        self.world = self.dp.world
        self.sync_bn = bool(sync_bn) and self.world > 1
        # forward binary GEMMs on tcgen05 kind::i8 (int8 +-1 operands, exact int32 accumulation):
        # half the operand bytes and twice the MMA rate of the bf16 path, bit-identical results
        self.forward_int8 = bool(forward_int8)
        self._shared = share_state_with
        # step counter and a version stamp of the weights (bumped by every update / load), shared
        # between engines that alias one state: cached inference preparations check the stamp
        self._state = share_state_with._state if share_state_with is not None else {"step": 0, "version": 0}
        self._eval_version = -1
        self.launches_per_step = 0
        # Two engine-owned streams: the dependent chain of a step runs on a high-priority one, the
        # weight-gradient GEMMs (nothing in the backward chain waits for them) on a low-priority
        # one, so that the tensor-pipe-bound wgrad of layer i overlaps the HBM-bound BatchNorm
        # backward of layer i-1.
        self.overlap = True
        lo_pri, hi_pri = torch.cuda.Stream.priority_range() if hasattr(torch.cuda.Stream, "priority_range") else (0, -1)
        self._main = torch.cuda.Stream(device=self.device, priority=hi_pri)
        self._side = torch.cuda.Stream(device=self.device, priority=lo_pri)
        self._side_events = [torch.cuda.Event() for _ in range(len(self.layers) + 1)]
        self._wgrad_events = [torch.cuda.Event() for _ in range(len(self.layers) + 1)]
        self._prep_fork, self._prep_join, self._prep_pending = torch.cuda.Event(), torch.cuda.Event(), False
        self._pending_reduce = []      # in-flight gradient-bucket all-reduces of the current step
        self.reduce_gradients = True   # False: keep gradients local (accumulation / compute-only timing)
        # first layer's backward inside layer 1's dgrad epilogue (ZKB_FUSE_FIRST_BWD=0: separate kernels)
        self.fuse_first_backward = (os.environ.get("ZKB_FUSE_FIRST_BWD", "1") != "0"
                                    and self._can_fuse_first_backward())
        # BatchNorm-backward sums of layer i-1 inside layer i's dgrad epilogue (conv layers)
        self.fuse_bn_backward_sums = os.environ.get("ZKB_FUSE_BN_BWD", "1") != "0"
        # BatchNorm batch statistics inside the forward GEMM epilogue (layers that do not pool)
        self.fuse_forward_stats = os.environ.get("ZKB_FUSE_FWD_STATS", "1") != "0"
        # 2x2 max-pool (+ statistics) inside the forward GEMM epilogue of pooling layers whose map
        # tiles into whole row pairs (ZKB_FUSE_POOL=0: separate zkb_maxpool2_i16)
        self.fuse_pool = self.forward_int8 and os.environ.get("ZKB_FUSE_POOL", "1") != "0"
        # inference: BatchNorm + sign as integer thresholds inside the GEMM epilogue
        self.fold_eval = self.forward_int8 and os.environ.get("ZKB_FOLD_EVAL", "1") != "0"
        self.watch = None      # None | True | set of entry-point names to time with CUDA events
        self.watched = []      # (name, layer tag, start event, end event)
        self._tag = ""
        self._validate()
        self._allocate()

    @property
    def step_count(self) -> int:
        return self._state["step"]

    @step_count.setter
    def step_count(self, v: int) -> None:
        self._state["step"] = int(v)

    def _touch(self) -> None:
        """The weights or moving statistics change: cached inference preparations are stale."""
        self._state["version"] += 1

    @property
    def _eval_ready(self) -> bool:
        return self._eval_version == self._state["version"]

    # ------------------------------------------------------------------ set-up
    def _validate(self):
        ls = self.layers
        if not ls or ls[0].kind != "conv" or ls[0].quant_input or ls[0].padding != "valid" or ls[0].pool:
            raise ValueError("first layer must be an unquantised-input 'valid' conv without pooling")
        if ls[0].ksize != 3 or ls[0].cin not in (1, 3):
            raise ValueError("first layer: 3x3 kernel over 1 or 3 input channels")
        for l in ls[1:]:
            if not l.quant_input:
                raise ValueError("only the first layer may skip the input quantizer")
            if l.kind == "conv" and (l.padding != "same" or l.ksize % 2 != 1):
                raise ValueError("binary convs must be odd-sized 'same' convs")
            if l.cin % 128 != 0:
                raise ValueError(f"binary layer input width {l.cin} must be a multiple of 128")
        if ls[-1].kind != "dense" or ls[-1].cout != self.num_classes:
            raise ValueError("last layer must be a dense layer with num_classes units")
        if ls[-1].pool:
            raise ValueError("pooling after a dense layer")

    def _allocate(self):
        dev, B = self.device, self.B
        H, W, C = self.input_shape
        f32, i16, bf16, u8 = torch.float32, torch.int16, torch.bfloat16, torch.uint8
        # ---- parameter arenas
        k_sizes, b_sizes = [], []
        for l in self.layers:
            k_sizes.append(l.ksize * l.ksize * l.cin * l.cout if l.kind == "conv" else l.cin * l.cout)
            b_sizes.append(self._cpad(l))
        self.k_off = np.concatenate([[0], np.cumsum([_round_up(s, 4) for s in k_sizes])]).astype(np.int64)
        self.b_off = np.concatenate([[0], np.cumsum(b_sizes)]).astype(np.int64)
        self.n_kernel = int(self.k_off[-1])
        self.n_beta = int(self.b_off[-1])
        n = self.n_kernel + self.n_beta
        # Arena order [all betas | kernel 0 | kernel 1 | ...]: the gradients that are ready LAST in the
        # backward pass (betas of every layer, kernels of the first layers) are one contiguous
        # prefix, those ready first (the dense kernels) a suffix -- the reverse-layer-order buckets of
        # the gradient all-reduce are plain slices.
        self.kbase = self.n_beta
        if self._shared is not None:
            o = self._shared
            if o.params.numel() != n or o.n_kernel != self.n_kernel or o.device != dev:
                raise ValueError("share_state_with: the other engine holds a different model or device")
            self.params, self.adam_m, self.adam_v = o.params, o.adam_m, o.adam_v
            self.moving_mean, self.moving_var = o.moving_mean, o.moving_var
        else:
            self.params = torch.zeros(n, dtype=f32, device=dev)
            self.adam_m = torch.zeros(n, dtype=f32, device=dev)
            self.adam_v = torch.zeros(n, dtype=f32, device=dev)
            self.moving_mean = torch.zeros(self.n_beta, dtype=f32, device=dev)
            self.moving_var = torch.ones(self.n_beta, dtype=f32, device=dev)
        self.grads = torch.zeros(n, dtype=f32, device=dev)
        # per-step accumulators (zeroed by one memset): fwd sums, bwd sums, loss, correct
        # (+ the fused first-layer path: 11 sums per channel from the dgrad epilogue, and the
        # Gram matrix of the input taps G[9][9] + S2[9])
        self.acc = torch.zeros(4 * self.n_beta + 2 + 11 * 128 + 90, dtype=torch.float64, device=dev)
        self.l0_acc = self.acc[4 * self.n_beta + 2: 4 * self.n_beta + 2 + 11 * 128]
        self.l0_gram = self.acc[4 * self.n_beta + 2 + 11 * 128:]
        self.correct = torch.zeros(1, dtype=torch.int32, device=dev)
        self.stat = torch.zeros(4 * self.n_beta, dtype=f32, device=dev)  # mean, rstd, mean_da, mean_dax
        self.x0 = torch.zeros(B, H, W, C, dtype=f32, device=dev)
        self.labels = torch.zeros(B, dtype=torch.int32, device=dev)
        self.L: List[_Layer] = []
        h, w = H, W
        for i, l in enumerate(self.layers):
            L = _Layer()
            L.spec, L.idx = l, i
            L.cp = self._cpad(l)
            ksz = k_sizes[i]
            ko = self.kbase + int(self.k_off[i])
            L.kernel = self.params[ko: ko + ksz]
            L.dkernel = self.grads[ko: ko + ksz]
            bo = int(self.b_off[i])
            L.beta = self.params[bo: bo + L.cp]
            L.dbeta = self.grads[bo: bo + L.cp]
            so = int(self.b_off[i])
            nb = self.n_beta
            # accumulators of one layer are adjacent, [sum | sumsq | sum_da | sum_dax]: under
            # synchronised BatchNorm the forward pair and the backward pair are ONE collective each
            ao = 4 * so
            L.sum, L.sumsq = self.acc[ao: ao + L.cp], self.acc[ao + L.cp: ao + 2 * L.cp]
            L.sum_da, L.sum_dax = self.acc[ao + 2 * L.cp: ao + 3 * L.cp], self.acc[ao + 3 * L.cp: ao + 4 * L.cp]
            L.fwd_sums, L.bwd_sums = self.acc[ao: ao + 2 * L.cp], self.acc[ao + 2 * L.cp: ao + 4 * L.cp]
            L.mean, L.rstd = self.stat[so: so + L.cp], self.stat[nb + so: nb + so + L.cp]
            L.mean_da, L.mean_dax = self.stat[2 * nb + so: 2 * nb + so + L.cp], self.stat[3 * nb + so: 3 * nb + so + L.cp]
            L.mm, L.mv = self.moving_mean[so: so + L.cp], self.moving_var[so: so + L.cp]
            if l.kind == "conv":
                L.in_hw = (h, w)
                oh, ow = (h - l.ksize + 1, w - l.ksize + 1) if l.padding == "valid" else (h, w)
                L.out_hw = (oh, ow)
                # the first layer's float map is never stored: its K=9 conv is recomputed inside
                # the fused BatchNorm kernels (zkb_conv_first_bn_* / _bwd_*)
                L.y = None if i == 0 else torch.zeros(B, oh, ow, L.cp, dtype=i16, device=dev)
                if l.pool:
                    ph, pw = oh // 2, ow // 2
                    L.p = torch.zeros(B, ph, pw, L.cp, dtype=i16, device=dev)
                    L.arg = torch.zeros(B, ph, pw, L.cp, dtype=u8, device=dev)
                    h, w = ph, pw
                else:
                    h, w = oh, ow
                L.bn_rows = B * h * w
                L.s = torch.zeros(B, h, w, L.cp, dtype=bf16, device=dev)
                L.s8 = torch.zeros(B, h, w, L.cp, dtype=torch.int8, device=dev) if self.forward_int8 else None
                # with the fused first-layer backward the gradient of the first sign map lives only
                # in tensor memory (1.4 GB of HBM at batch 4096 never allocated)
                L.ds = None if (i == 0 and self.fuse_first_backward) else torch.zeros(B, h, w, L.cp, dtype=f32, device=dev)
                if i == 0:
                    L.wq = torch.zeros(l.ksize * l.ksize * l.cin, l.cout, dtype=f32, device=dev)
                else:
                    L.w_fwd = torch.zeros(l.cout, l.ksize * l.ksize * l.cin, dtype=bf16, device=dev)
                    L.w_fwd8 = torch.zeros(l.cout, l.ksize * l.ksize * l.cin, dtype=torch.int8, device=dev) if self.forward_int8 else None
                    L.w_dgrad = torch.zeros(l.cin, l.ksize * l.ksize * l.cout, dtype=bf16, device=dev)
                    L.dy_hi = torch.zeros(B, oh, ow, L.cp, dtype=bf16, device=dev)
                    L.dy_lo = torch.zeros(B, oh, ow, L.cp, dtype=bf16, device=dev)
            else:
                L.y = torch.zeros(B, L.cp, dtype=i16, device=dev)
                L.bn_rows = B
                last = i == len(self.layers) - 1
                if last:
                    L.logits = torch.zeros(B, L.cp, dtype=f32, device=dev)
                    L.ds = torch.zeros(B, L.cp, dtype=f32, device=dev)  # dlogits
                else:
                    L.s = torch.zeros(B, L.cp, dtype=bf16, device=dev)
                    L.s8 = torch.zeros(B, L.cp, dtype=torch.int8, device=dev) if self.forward_int8 else None
                    L.ds = torch.zeros(B, L.cp, dtype=f32, device=dev)
                L.w_fwd = torch.zeros(L.cp, l.cin, dtype=bf16, device=dev)
                L.w_fwd8 = torch.zeros(L.cp, l.cin, dtype=torch.int8, device=dev) if self.forward_int8 else None
                L.w_dgrad = torch.zeros(l.cin, L.cp, dtype=bf16, device=dev)
                L.dy_hi = torch.zeros(B, L.cp, dtype=bf16, device=dev)
                L.dy_lo = torch.zeros(B, L.cp, dtype=bf16, device=dev)
            self.L.append(L)
        self.grad_buckets = self._make_grad_buckets()

    def _make_grad_buckets(self):
        return gradient_buckets(self.k_off, self.kbase)

    @staticmethod
    def _cpad(l: LayerSpec) -> int:
        return l.cout if l.cout % 64 == 0 else _round_up(l.cout, NPAD)

    # ------------------------------------------------------------------ parameters in / out
    def load_params(self, params: Dict[str, np.ndarray], moving: Optional[Dict[str, np.ndarray]] = None):
        """Inject weights in Keras layout: ``kernel{i}`` HWIO / [K, N], ``beta{i}`` [C]."""
        self._touch()
        for L in self.L:
            k = torch.as_tensor(np.ascontiguousarray(params[f"kernel{L.idx}"]), dtype=torch.float32)
            if k.numel() != L.kernel.numel():
                raise ValueError(f"kernel{L.idx}: expected {L.kernel.numel()} values, got {k.numel()}")
            L.kernel.copy_(k.reshape(-1))
            b = torch.as_tensor(params[f"beta{L.idx}"], dtype=torch.float32)
            L.beta.zero_()
            L.beta[: L.spec.cout].copy_(b)
            if moving is not None:
                L.mm.zero_()
                L.mv.fill_(1.0)
                L.mm[: L.spec.cout].copy_(torch.as_tensor(moving[f"moving_mean{L.idx}"], dtype=torch.float32))
                L.mv[: L.spec.cout].copy_(torch.as_tensor(moving[f"moving_var{L.idx}"], dtype=torch.float32))

    def init_glorot(self, seed: int = 0):
        """Keras default (Glorot uniform) initialisation from ``numpy.random.default_rng(seed)``."""
        rng = np.random.default_rng(seed)
        params = {}
        for i, l in enumerate(self.layers):
            if l.kind == "conv":
                shape = (l.ksize, l.ksize, l.cin, l.cout)
                fi, fo = l.ksize * l.ksize * l.cin, l.ksize * l.ksize * l.cout
            else:
                shape, fi, fo = (l.cin, l.cout), l.cin, l.cout
            lim = math.sqrt(6.0 / (fi + fo))
            params[f"kernel{i}"] = rng.uniform(-lim, lim, size=shape).astype(np.float32)
            params[f"beta{i}"] = np.zeros(l.cout, dtype=np.float32)
        self.load_params(params)
        return params

    def _shape_of(self, L) -> Tuple[int, ...]:
        l = L.spec
        return (l.ksize, l.ksize, l.cin, l.cout) if l.kind == "conv" else (l.cin, l.cout)

    def get_params(self) -> Dict[str, np.ndarray]:
        out = {}
        for L in self.L:
            out[f"kernel{L.idx}"] = L.kernel.detach().cpu().numpy().reshape(self._shape_of(L)).copy()
            out[f"beta{L.idx}"] = L.beta[: L.spec.cout].detach().cpu().numpy().copy()
        return out

    def get_grads(self) -> Dict[str, np.ndarray]:
        out = {}
        for L in self.L:
            out[f"kernel{L.idx}"] = L.dkernel.detach().cpu().numpy().reshape(self._shape_of(L)).copy()
            out[f"beta{L.idx}"] = L.dbeta[: L.spec.cout].detach().cpu().numpy().copy()
        return out

    def get_adam_state(self) -> Tuple[Dict[str, np.ndarray], Dict[str, np.ndarray]]:
        """First and second Adam moments in the same keyed layout as :meth:`get_params`."""
        m, v = {}, {}
        for L in self.L:
            k0, ksz = self.kbase + int(self.k_off[L.idx]), L.kernel.numel()
            b0 = int(self.b_off[L.idx])
            for dst, arena in ((m, self.adam_m), (v, self.adam_v)):
                dst[f"kernel{L.idx}"] = arena[k0: k0 + ksz].detach().cpu().numpy().reshape(self._shape_of(L)).copy()
                dst[f"beta{L.idx}"] = arena[b0: b0 + L.spec.cout].detach().cpu().numpy().copy()
        return m, v

    # ------------------------------------------------------------------ checkpoints
    def save_checkpoint(self, path: str) -> None:
        """Weights (Keras layout: ``kernel{i}`` HWIO / [K, N], ``beta{i}``), BatchNorm moving
        statistics, Adam moments and the step counter in one ``.npz``."""
        blob = dict(self.get_params())
        blob.update(self.get_moving())
        m, v = self.get_adam_state()
        blob.update({f"adam_m/{k}": a for k, a in m.items()})
        blob.update({f"adam_v/{k}": a for k, a in v.items()})
        blob["step_count"] = np.asarray(self.step_count, dtype=np.int64)
        np.savez(path, **blob)

    def load_checkpoint(self, path: str) -> None:
        with np.load(path) as z:
            blob = {k: z[k] for k in z.files}
        self.load_params(blob, blob if "moving_mean0" in blob else None)
        for L in self.L:
            k0, ksz = self.kbase + int(self.k_off[L.idx]), L.kernel.numel()
            b0 = int(self.b_off[L.idx])
            for tag, arena in (("adam_m", self.adam_m), ("adam_v", self.adam_v)):
                if f"{tag}/kernel{L.idx}" in blob:
                    arena[k0: k0 + ksz].copy_(torch.as_tensor(blob[f"{tag}/kernel{L.idx}"], dtype=torch.float32).reshape(-1))
                    arena[b0: b0 + L.spec.cout].copy_(torch.as_tensor(blob[f"{tag}/beta{L.idx}"], dtype=torch.float32))
        self.step_count = int(blob.get("step_count", 0))

    def average_moving_stats(self) -> None:
        """Per-replica BatchNorm: every rank has folded its own batch statistics into its moving
        mean / variance, so what a validation pass or a checkpoint sees would depend on the rank.
        Replace them by their mean over the replicas.  Nothing to do on one
        GPU or under synchronised BatchNorm (identical on every rank by construction)."""
        if self.world == 1 or self.sync_bn:
            return
        for t in (self.moving_mean, self.moving_var):
            self._all_reduce(t)
            t.mul_(1.0 / self.world)
        self._touch()

    def get_moving(self) -> Dict[str, np.ndarray]:
        out = {}
        for L in self.L:
            out[f"moving_mean{L.idx}"] = L.mm[: L.spec.cout].detach().cpu().numpy().copy()
            out[f"moving_var{L.idx}"] = L.mv[: L.spec.cout].detach().cpu().numpy().copy()
        return out

    # ------------------------------------------------------------------ kernels
    def _call(self, name, *args, on=None):
        """Launch entry point ``name`` on the step's main stream, or on torch stream ``on``."""
        self._n_launch += 1
        handle = self._stream if on is None else on.cuda_stream
        watch = self.watch
        if watch is not None and (watch is True or name in watch):
            # CUDA events on the launching stream around this one launch (bench.py roofline /
            # per-kernel breakdown); off by default.
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record(on) if on is not None else e0.record()
            zn.call(name, *args, handle)
            e1.record(on) if on is not None else e1.record()
            self.watched.append((name, self._tag, e0, e1))
            return
        zn.call(name, *args, handle)

    def _begin(self):
        self._stream = torch.cuda.current_stream(self.device).cuda_stream
        self._n_launch = 0

    def preprocess(self, images_u8: torch.Tensor, labels_i64: Optional[torch.Tensor], index: Optional[torch.Tensor],
                   training: bool, pad_size: int = 0, aug: Optional[torch.Tensor] = None, seed: int = 0,
                   step: int = 0, binarize: bool = False, onehot: Optional[torch.Tensor] = None, batch: Optional[int] = None,
                   step_dev: Optional[torch.Tensor] = None):
        """Fused batch map of ``Preprocessing.__call__`` for the example's PadCropAndFlip
        (zookeeper/tf/preprocessing.py:59-75, examples/larq_experiment.py:24-37)."""
        if not hasattr(self, "_stream"):
            self._begin()
        B = self.B if batch is None else batch
        _, IH, IW, C = images_u8.shape
        H, W, _ = self.input_shape
        self._call("zkb_preprocess_u8", zn.ptr(images_u8), zn.ptr(labels_i64), zn.ptr(index), zn.ptr(aug),
                   zn.ptr(self.x0), zn.ptr(self.labels) if labels_i64 is not None else 0, zn.ptr(onehot),
                   B, IH, IW, C, H, W, pad_size, int(training), int(binarize), self.num_classes, seed, step,
                   zn.ptr(step_dev))

    def _all_reduce(self, t: torch.Tensor):
        self.dp.all_reduce_sum_(t)

    def _bn_forward(self, L, x, is_i16: bool, rows: int, stats_done: bool, training: bool = True):
        cp = L.cp
        if training:
            if not stats_done:
                self._call("zkb_bn_stats", zn.ptr(x), int(is_i16), rows, cp, zn.ptr(L.sum), zn.ptr(L.sumsq))
            count = float(rows)
            if self.sync_bn:
                self._pending_bn_sync(L)
                count *= self.world
            bessel = 1 if L.spec.kind == "conv" else 0
            self._call("zkb_bn_finalize", zn.ptr(L.sum), zn.ptr(L.sumsq), count, BN_EPS, BN_MOMENTUM, bessel, cp,
                       zn.ptr(L.mean), zn.ptr(L.rstd), zn.ptr(L.mm), zn.ptr(L.mv))
            mean, rstd = L.mean, L.rstd
        else:
            mean, rstd = L.mm, L.eval_rstd
        last = L.idx == len(self.L) - 1
        self._call("zkb_bn_apply_sign", zn.ptr(x), int(is_i16), rows, cp, zn.ptr(mean), zn.ptr(rstd), zn.ptr(L.beta),
                   0 if last else zn.ptr(L.s), zn.ptr(L.logits) if last else 0,
                   0 if last else zn.ptr(getattr(L, "s8", None)))

    def _pending_bn_sync(self, L):
        self._all_reduce(L.fwd_sums)     # [sum | sumsq] of this layer in one collective

    def forward(self, training: bool = True):
        """Forward pass over ``self.x0`` (filled by :meth:`preprocess`)."""
        B = self.B
        prev = None
        for L in self.L:
            l = L.spec
            self._tag = f"fwd{L.idx}"
            if L.idx >= 1 and self._prep_pending:       # binary-layer weights prepared on the side stream
                torch.cuda.current_stream(self.device).wait_event(self._prep_join)
                self._prep_pending = False
            if l.kind == "conv":
                oh, ow = L.out_hw
                if L.idx == 0:
                    self._first_layer_forward(L, training)
                else:
                    ih, iw = L.in_hw
                    fused_stats = self.forward_int8 and self.fuse_forward_stats and training and not l.pool
                    if l.pool and training and self._pool_in_epilogue(L):
                        # conv + pool + statistics in one launch; the full-resolution map is never stored
                        self._call("zkb_bconv2d_fwd_i8_pool", zn.ptr(prev.s8), zn.ptr(L.w_fwd8), zn.ptr(L.p), zn.ptr(L.arg),
                                   B, ih, iw, l.cin, l.cout, l.ksize, zn.ptr(L.sum), zn.ptr(L.sumsq))
                        self._bn_forward(L, L.p, True, L.bn_rows, True, training)
                        prev = L
                        continue
                    if fused_stats:
                        self._call("zkb_bconv2d_fwd_i8_stats", zn.ptr(prev.s8), zn.ptr(L.w_fwd8), zn.ptr(L.y), B, ih, iw,
                                   l.cin, l.cout, l.ksize, zn.ptr(L.sum), zn.ptr(L.sumsq))
                    elif self.forward_int8:
                        self._call("zkb_bconv2d_fwd_i8", zn.ptr(prev.s8), zn.ptr(L.w_fwd8), zn.ptr(L.y), B, ih, iw, l.cin, l.cout, l.ksize)
                    else:
                        self._call("zkb_bconv2d_fwd", zn.ptr(prev.s), zn.ptr(L.w_fwd), zn.ptr(L.y), B, ih, iw, l.cin, l.cout, l.ksize)
                    if l.pool:
                        self._call("zkb_maxpool2_i16", zn.ptr(L.y), B, oh, ow, L.cp, zn.ptr(L.p), zn.ptr(L.arg),
                                   zn.ptr(L.sum), zn.ptr(L.sumsq))
                        self._bn_forward(L, L.p, True, L.bn_rows, True, training)
                    else:
                        self._bn_forward(L, L.y, True, L.bn_rows, fused_stats, training)
            else:
                fused_stats = self.forward_int8 and self.fuse_forward_stats and training
                if fused_stats:
                    self._call("zkb_bdense_fwd_i8_stats", zn.ptr(prev.s8), zn.ptr(L.w_fwd8), zn.ptr(L.y), L.cp, B, l.cin, L.cp,
                               zn.ptr(L.sum), zn.ptr(L.sumsq))
                elif self.forward_int8:
                    self._call("zkb_bdense_fwd_i8", zn.ptr(prev.s8), zn.ptr(L.w_fwd8), zn.ptr(L.y), L.cp, B, l.cin, L.cp)
                else:
                    self._call("zkb_bdense_fwd", zn.ptr(prev.s), zn.ptr(L.w_fwd), zn.ptr(L.y), L.cp, B, l.cin, L.cp)
                self._bn_forward(L, L.y, True, B, fused_stats, training)
            prev = L

    def _first_layer_forward(self, L, training: bool):
        l = L.spec
        H, W, C = self.input_shape
        geom = (self.B, H, W, C, l.cout, l.ksize)
        if training:
            if self.fuse_first_backward:
                # batch statistics of the first conv from the Gram matrix of its 9 input taps (the
                # backward needs the same 90 numbers): one pass over the 12.8 MB of images
                self._call("zkb_conv_first_gram", zn.ptr(self.x0), self.B, H, W, self.l0_gram.data_ptr())
                self._call("zkb_conv_first_stats_from_gram", self.l0_gram.data_ptr(), zn.ptr(L.wq), l.cout,
                           zn.ptr(L.sum), zn.ptr(L.sumsq))
            else:
                self._call("zkb_conv_first_bn_stats", zn.ptr(self.x0), zn.ptr(L.wq), *geom, zn.ptr(L.sum), zn.ptr(L.sumsq))
            count = float(L.bn_rows)
            if self.sync_bn:
                self._pending_bn_sync(L)
                count *= self.world
            self._call("zkb_bn_finalize", zn.ptr(L.sum), zn.ptr(L.sumsq), count, BN_EPS, BN_MOMENTUM, 1, L.cp,
                       zn.ptr(L.mean), zn.ptr(L.rstd), zn.ptr(L.mm), zn.ptr(L.mv))
            mean, rstd = L.mean, L.rstd
        else:
            mean, rstd = L.mm, L.eval_rstd
        self._call("zkb_conv_first_bn_sign", zn.ptr(self.x0), zn.ptr(L.wq), *geom, zn.ptr(mean), zn.ptr(rstd),
                   zn.ptr(L.beta), zn.ptr(L.s), zn.ptr(L.s8))

    def _pool_in_epilogue(self, L) -> bool:
        """zkb_bconv2d_fwd_i8_pool covers this pooling layer (even map whose row pairs tile 256 MMA
        columns, channel count a multiple of 128)."""
        if not self.fuse_pool or L.spec.kind != "conv" or not L.spec.pool or L.cp != L.spec.cout:
            return False
        oh, ow = L.out_hw
        lib = zn.load()
        return bool(lib.zkb_bconv2d_fwd_i8_pool_ok(int(oh), int(ow), int(L.spec.cout)))

    def _can_fuse_first_backward(self) -> bool:
        """zkb_bconv2d_dgrad_first_bwd is built for: 3x3 'valid' conv of one input channel into 128
        channels (no pooling), followed by a 3x3 'same' conv -- the MNIST topology of the example."""
        if len(self.layers) < 2:
            return False
        a, b = self.layers[0], self.layers[1]
        return (a.kind == "conv" and a.ksize == 3 and a.padding == "valid" and a.cin == 1 and a.cout == 128
                and not a.pool and b.kind == "conv" and b.ksize == 3 and b.padding == "same")

    def _sums_come_from_dgrad(self, i: int) -> bool:
        """Layer i's BatchNorm-backward sums are produced by the dgrad of layer i+1 (a conv whose
        input is layer i's sign map, channel count a multiple of 128)."""
        if not self.fuse_bn_backward_sums or i < 1 or i + 1 >= len(self.layers):
            return False
        nxt = self.layers[i + 1]
        return nxt.kind == "conv" and self.layers[i].kind == "conv" and self.L[i].cp % 128 == 0

    def _fused_first_layer_backward(self, L1):
        """Layer 1's dgrad GEMM whose epilogue is the whole backward of layer 0 (BatchNorm sums,
        ste mask, weight-gradient sums), then a one-block finalize."""
        L0 = self.L[0]
        H, W, _ = self.input_shape
        ih, iw = L1.in_hw
        l = L1.spec
        self._call("zkb_bconv2d_dgrad_first_bwd", zn.ptr(L1.dy_hi), zn.ptr(L1.dy_lo), zn.ptr(L1.w_dgrad),
                   self.B, ih, iw, l.cin, l.cout, l.ksize, zn.ptr(self.x0), H, W, zn.ptr(L0.wq),
                   zn.ptr(L0.mean), zn.ptr(L0.rstd), zn.ptr(L0.beta), self.l0_acc.data_ptr())
        count = float(L0.bn_rows)
        if self.sync_bn:
            self._all_reduce(self.l0_acc[:256])      # sum da, sum da*xhat; the tap sums stay local
            count *= self.world
        self._tag = "bwd0"
        self._call("zkb_conv_first_bwd_finalize", self.l0_acc.data_ptr(), self.l0_gram.data_ptr(), zn.ptr(L0.wq),
                   zn.ptr(L0.mean), zn.ptr(L0.rstd), count, (1.0 / self.world) if self.sync_bn else 1.0,
                   zn.ptr(L0.dbeta), zn.ptr(L0.dkernel), L0.spec.cout)

    def _first_layer_backward(self, L):
        l = L.spec
        H, W, C = self.input_shape
        geom = (self.B, H, W, C, l.cout, l.ksize)
        self._call("zkb_conv_first_bwd_reduce", zn.ptr(self.x0), zn.ptr(L.wq), zn.ptr(L.ds), *geom, zn.ptr(L.mean),
                   zn.ptr(L.rstd), zn.ptr(L.beta), zn.ptr(L.sum_da), zn.ptr(L.sum_dax))
        count = float(L.bn_rows)
        if self.sync_bn:
            self._all_reduce(L.bwd_sums)
            count *= self.world
        self._call("zkb_bn_bwd_finalize", zn.ptr(L.sum_da), zn.ptr(L.sum_dax), count,
                   (1.0 / self.world) if self.sync_bn else 1.0, L.cp, zn.ptr(L.dbeta),
                   zn.ptr(L.mean_da), zn.ptr(L.mean_dax))
        self._call("zkb_conv_first_bwd_wgrad", zn.ptr(self.x0), zn.ptr(L.wq), zn.ptr(L.ds), *geom, zn.ptr(L.mean),
                   zn.ptr(L.rstd), zn.ptr(L.beta), zn.ptr(L.mean_da), zn.ptr(L.mean_dax), zn.ptr(L.dkernel))

    def prepare_weights(self, side=None):
        """sign(W) in the operand layouts of the GEMMs.  With ``side`` (a second stream) the binary
        layers' weights are prepared there, next to the first layer's forward on the current
        stream; :meth:`forward` waits for them before the first binary GEMM."""
        self._tag = "prep"
        main = torch.cuda.current_stream(self.device)
        if side is not None:
            self._prep_fork.record(main)     # after the previous step's Adam
            side.wait_event(self._prep_fork)
        for L in self.L:
            l = L.spec
            if L.idx == 0:
                self._call("zkb_sign_f32", zn.ptr(L.kernel), zn.ptr(L.wq), L.kernel.numel())
            elif l.kind == "conv":
                self._call("zkb_prep_conv_weights", zn.ptr(L.kernel), l.ksize, l.cin, l.cout, zn.ptr(L.w_fwd), zn.ptr(L.w_dgrad),
                           zn.ptr(L.w_fwd8), on=side)
            else:
                self._call("zkb_prep_dense_weights", zn.ptr(L.kernel), l.cin, l.cout, L.cp, zn.ptr(L.w_fwd), zn.ptr(L.w_dgrad),
                           zn.ptr(L.w_fwd8), on=side)
        if side is not None:
            self._prep_join.record(side)
            self._prep_pending = True

    def backward(self, side=None, overlap_reduce: bool = False):
        """Backward pass.  With ``side`` (a low-priority stream) the wgrad GEMMs are issued there:
        each waits for its layer's dy and is joined before the gradients are consumed.  With
        ``overlap_reduce`` (data-parallel runs) the gradient buckets whose last weight-gradient GEMM
        has been enqueued are all-reduced right behind it, next to the rest of the backward pass;
        :meth:`finish_gradient_reduce` issues the last bucket and joins them all."""
        B = self.B
        n = len(self.L)
        main = torch.cuda.current_stream(self.device)
        self._pending_reduce = []
        early = ({b[0]: b for b in self.grad_buckets[:-1]}
                 if (overlap_reduce and side is not None and self.world > 1 and self.reduce_gradients) else {})
        for i in range(n - 1, -1, -1):
            L, l = self.L[i], self.L[i].spec
            self._tag = f"bwd{i}"
            if i == 0:
                if not self.fuse_first_backward:
                    self._first_layer_backward(L)
                continue
            last = i == n - 1
            x = L.p if (l.kind == "conv" and l.pool) else L.y
            is_i16 = 1
            rows = L.bn_rows
            ste = 0 if last else 1
            if not self._sums_come_from_dgrad(i):
                self._call("zkb_bn_bwd_reduce", zn.ptr(L.ds), zn.ptr(x), is_i16, rows, L.cp, zn.ptr(L.mean), zn.ptr(L.rstd),
                           zn.ptr(L.beta), ste, zn.ptr(L.sum_da), zn.ptr(L.sum_dax))
            count = float(rows)
            if self.sync_bn:
                self._all_reduce(L.bwd_sums)     # [sum_da | sum_dax] in one collective
                count *= self.world
            if l.kind == "conv" and l.pool:
                mode, of32, ohi, olo, arg = 2, 0, zn.ptr(L.dy_hi), zn.ptr(L.dy_lo), zn.ptr(L.arg)
                H, W = L.out_hw
            else:
                mode, of32, ohi, olo, arg, H, W = 1, 0, zn.ptr(L.dy_hi), zn.ptr(L.dy_lo), 0, 0, 0
            dbeta_scale = (1.0 / self.world) if self.sync_bn else 1.0
            self._call("zkb_bn_bwd_finalize", zn.ptr(L.sum_da), zn.ptr(L.sum_dax), count, dbeta_scale, L.cp,
                       zn.ptr(L.dbeta), zn.ptr(L.mean_da), zn.ptr(L.mean_dax))
            self._call("zkb_bn_bwd_apply", zn.ptr(L.ds), zn.ptr(x), is_i16, rows, L.cp, zn.ptr(L.mean), zn.ptr(L.rstd),
                       zn.ptr(L.beta), zn.ptr(L.mean_da), zn.ptr(L.mean_dax), ste, mode, of32, ohi, olo, arg, H, W)
            prev = self.L[i - 1]
            if side is not None:
                ev = self._side_events[i]
                ev.record(main)          # dy_i is complete (and the gradient arena was zeroed) here
                side.wait_event(ev)
            if l.kind == "conv":
                ih, iw = L.in_hw
                # dgrad first: it is on the critical path (the next BatchNorm backward needs it)
                if i == 1 and self.fuse_first_backward:
                    self._fused_first_layer_backward(L)
                    self._tag = f"bwd{i}"
                elif self._sums_come_from_dgrad(i - 1):
                    px = prev.p if prev.spec.pool else prev.y
                    self._call("zkb_bconv2d_dgrad_bn", zn.ptr(L.dy_hi), zn.ptr(L.dy_lo), zn.ptr(L.w_dgrad), zn.ptr(prev.ds),
                               B, ih, iw, l.cin, l.cout, l.ksize, zn.ptr(px), zn.ptr(prev.mean), zn.ptr(prev.rstd),
                               zn.ptr(prev.beta), zn.ptr(prev.sum_da), zn.ptr(prev.sum_dax))
                else:
                    self._call("zkb_bconv2d_dgrad", zn.ptr(L.dy_hi), zn.ptr(L.dy_lo), zn.ptr(L.w_dgrad), zn.ptr(prev.ds),
                               B, ih, iw, l.cin, l.cout, l.ksize)
                self._call("zkb_bconv2d_wgrad", zn.ptr(prev.s), zn.ptr(L.dy_hi), zn.ptr(L.dy_lo), zn.ptr(L.dkernel),
                           B, ih, iw, l.cin, l.cout, l.ksize, on=side)
            else:
                self._call("zkb_bdense_dgrad", zn.ptr(L.dy_hi), zn.ptr(L.dy_lo), zn.ptr(L.w_dgrad), zn.ptr(prev.ds),
                           B, l.cin, L.cp)
                self._call("zkb_bdense_wgrad", zn.ptr(prev.s), zn.ptr(L.dy_hi), zn.ptr(L.dy_lo), zn.ptr(L.dkernel),
                           l.cout, B, l.cin, L.cp, l.cout, on=side)
            if i in early:
                # every kernel gradient of this bucket has been enqueued on the side stream: the
                # collective is ordered behind them there and runs next to the remaining backward
                _, lo, hi = early[i]
                with torch.cuda.stream(side):
                    self._pending_reduce.append(self.dp.all_reduce_sum_async_(self.grads[lo:hi]))
        if side is not None:
            main.wait_stream(side)       # every dW is in the arena before all-reduce / Adam

    def finish_gradient_reduce(self):
        """Sum the gradients over the replicas: the buckets :meth:`backward` has not started yet (all
        of them without overlap -- then the arena goes out as one collective), then wait for all."""
        if self.world == 1 or not self.reduce_gradients:
            return
        if not self._pending_reduce:
            self._all_reduce(self.grads)
            return
        _, lo, hi = self.grad_buckets[-1]
        self._pending_reduce.append(self.dp.all_reduce_sum_async_(self.grads[lo:hi]))
        for work in self._pending_reduce:
            if work is not None:
                work.wait()          # orders the current stream behind the collective; no host block
        self._pending_reduce = []

    def loss_and_grad(self):
        L = self.L[-1]
        nb4 = 4 * self.n_beta
        self._call("zkb_softmax_ce", zn.ptr(L.logits), L.cp, zn.ptr(self.labels), self.B, self.num_classes,
                   1.0 / (self.B * self.world), zn.ptr(L.ds), self.acc[nb4:].data_ptr(), zn.ptr(self.correct))

    def _lr_t(self, t: int) -> float:
        # bias-corrected step size in fp32, as TF's ApplyAdam computes it from the fp32 beta powers
        f = np.float32
        return float(f(self.lr) * np.sqrt(f(1) - f(self.b2) ** f(t), dtype=f) / (f(1) - f(self.b1) ** f(t)))

    def optimizer_step(self, lr_dev: Optional[torch.Tensor] = None):
        """Adam + clip.  ``lr_dev``: read the step size from device memory (captured graphs)."""
        self._touch()
        if lr_dev is None:
            self.step_count += 1
        lr_t = self._lr_t(max(self.step_count, 1))
        kb = self.kbase
        self._call("zkb_adam_clip", self.params[kb:].data_ptr(), self.grads[kb:].data_ptr(), self.adam_m[kb:].data_ptr(),
                   self.adam_v[kb:].data_ptr(), self.n_kernel, lr_t, self.b1, self.b2, self.eps, 1.0, 1, zn.ptr(lr_dev))
        self._call("zkb_adam_clip", zn.ptr(self.params), zn.ptr(self.grads), zn.ptr(self.adam_m), zn.ptr(self.adam_v),
                   self.n_beta, lr_t, self.b1, self.b2, self.eps, 1.0, 0, zn.ptr(lr_dev))

    # ------------------------------------------------------------------ public steps
    def train_step(self, images_u8: torch.Tensor, labels_i64: torch.Tensor, index: Optional[torch.Tensor] = None,
                   pad_size: int = 0, aug: Optional[torch.Tensor] = None, seed: int = 0, apply_update: bool = True,
                   preprocessed: bool = False):
        """preprocess -> forward -> softmax-CE -> backward -> (all-reduce) -> Adam + clip.

        Everything is enqueued on the current CUDA stream; nothing synchronises with the host.
        Read :meth:`last_loss` / :meth:`last_accuracy` afterwards (that read synchronises)."""
        self._touch()
        # per-launch timing of everything (watch is True) wants a serial schedule
        use_side = self.overlap and self.watch is not True
        caller = torch.cuda.current_stream(self.device)
        if use_side:
            self._main.wait_stream(caller)
        with torch.cuda.stream(self._main if use_side else caller):
            self._begin()
            self.acc.zero_()
            self.correct.zero_()
            self.grads.zero_()
            if not preprocessed:
                self.preprocess(images_u8, labels_i64, index, True, pad_size, aug, seed, self.step_count)
            self.prepare_weights(self._side if use_side else None)
            self.forward(training=True)
            self.loss_and_grad()
            self.backward(self._side if use_side else None, overlap_reduce=True)
            self.finish_gradient_reduce()
            if apply_update:
                self.optimizer_step()
        if use_side:
            caller.wait_stream(self._main)
        self.launches_per_step = self._n_launch

    # ------------------------------------------------------------------ CUDA-graph replay
    def capture_train_step(self, images_u8: torch.Tensor, labels_i64: torch.Tensor, pad_size: int, seed: int = 0) -> None:
        """Capture one full train step (both streams, the all-reduce included) into a CUDA graph.

        The ~90 launches of a step cost more host time than GPU time at small batches; a replay
        is one launch.  What changes between steps lives in device memory and is refreshed before
        each replay: the sample indices (``self.index_buf``), the step counter that keys the
        augmentation RNG and Adam's bias-corrected step size.  ``images_u8`` / ``labels_i64`` are
        the resident dataset pool and must stay alive and in place.  Run at least one ordinary
        :meth:`train_step` first (first-use set-up must not happen inside a capture)."""
        dev = self.device
        self.index_buf = torch.zeros(self.B, dtype=torch.int32, device=dev)
        self._dyn_step = torch.zeros(1, dtype=torch.int64, device=dev)
        self._dyn_lr = torch.zeros(1, dtype=torch.float32, device=dev)
        self._dyn_host = torch.zeros(2, dtype=torch.float64).pin_memory()   # [step, lr_t] staging
        self._graph_pool = (images_u8, labels_i64)
        graph = torch.cuda.CUDAGraph()
        torch.cuda.synchronize(dev)
        saved_watch, self.watch = self.watch, None
        try:
            with torch.cuda.graph(graph, stream=self._main):
                self._begin()
                self.acc.zero_()
                self.correct.zero_()
                self.grads.zero_()
                self.preprocess(images_u8, labels_i64, self.index_buf, True, pad_size, None, seed, 0,
                                step_dev=self._dyn_step)
                self.prepare_weights(self._side if self.overlap else None)
                self.forward(training=True)
                self.loss_and_grad()
                self.backward(self._side if self.overlap else None)   # (one collective inside a capture)
                self.finish_gradient_reduce()
                self.optimizer_step(lr_dev=self._dyn_lr)
        finally:
            self.watch = saved_watch
        self.launches_per_step = self._n_launch
        self._graph = graph

    def train_step_graph(self, index: torch.Tensor) -> None:
        """Replay the captured step on the batch ``index`` (int32 device tensor of sample ids)."""
        caller = torch.cuda.current_stream(self.device)
        self._touch()
        self.step_count += 1
        self.index_buf.copy_(index, non_blocking=True)
        self._dyn_step.fill_(self.step_count - 1)
        self._dyn_lr.fill_(self._lr_t(self.step_count))
        self._main.wait_stream(caller)
        with torch.cuda.stream(self._main):
            self._graph.replay()
        caller.wait_stream(self._main)

    def eval_step(self, images_u8: torch.Tensor, labels_i64: torch.Tensor, index: Optional[torch.Tensor] = None):
        """Validation pass (larq_experiment.py:133-139): centre crop/pad, moving statistics."""
        self._begin()
        self.acc.zero_()
        self.correct.zero_()
        if not self._eval_ready:
            # weights and moving statistics only change in training steps: sign(W) layouts, the
            # inference rstd and the folded thresholds are prepared once per run of eval steps
            for L in self.L:
                if getattr(L, "eval_rstd", None) is None:
                    L.eval_rstd = torch.zeros_like(L.mv)
                self._call("zkb_bn_eval_rstd", zn.ptr(L.mv), BN_EPS, L.cp, zn.ptr(L.eval_rstd))
            self.prepare_weights()
            if self.fold_eval:
                for L in self.L[1:-1]:
                    if not hasattr(L, "thr"):
                        L.thr = torch.zeros(L.cp, dtype=torch.int32, device=self.device)
                    self._call("zkb_bn_sign_thresholds", zn.ptr(L.mm), zn.ptr(L.eval_rstd), zn.ptr(L.beta), L.cp,
                               zn.ptr(L.thr))
            self._eval_version = self._state["version"]
        self.preprocess(images_u8, labels_i64, index, False)
        if self.fold_eval:
            self._forward_eval_folded()
        else:
            self.forward(training=False)
        L = self.L[-1]
        self._call("zkb_softmax_ce", zn.ptr(L.logits), L.cp, zn.ptr(self.labels), self.B, self.num_classes,
                   1.0 / self.B, 0, self.acc[4 * self.n_beta:].data_ptr(), zn.ptr(self.correct))

    def _forward_eval_folded(self):
        """Inference forward with BatchNorm + sign folded into one integer threshold per channel
        (SURVEY.md 8f.1): every hidden binary layer is ONE int8 GEMM whose epilogue compares the
        int32 accumulator with the threshold and stores int8 signs -- no int16 map, no BatchNorm
        pass; pooling layers take the maximum of the four signs (equal to the sign of the
        normalised maximum because rstd > 0).  Bit-identical to ``forward(training=False)``."""
        B = self.B
        prev = None
        last = len(self.L) - 1
        for L in self.L:
            l = L.spec
            self._tag = f"fwd{L.idx}"
            if L.idx == 0:
                H, W, C = self.input_shape
                self._call("zkb_conv_first_bn_sign", zn.ptr(self.x0), zn.ptr(L.wq), B, H, W, C, l.cout, l.ksize,
                           zn.ptr(L.mm), zn.ptr(L.eval_rstd), zn.ptr(L.beta), 0, zn.ptr(L.s8))
            elif L.idx == last:
                self._call("zkb_bdense_fwd_i8", zn.ptr(prev.s8), zn.ptr(L.w_fwd8), zn.ptr(L.y), L.cp, B, l.cin, L.cp)
                self._call("zkb_bn_apply_sign", zn.ptr(L.y), 1, B, L.cp, zn.ptr(L.mm), zn.ptr(L.eval_rstd), zn.ptr(L.beta),
                           0, zn.ptr(L.logits), 0)
            else:
                if l.kind == "conv":
                    ih, iw = L.in_hw
                    oh, ow = L.out_hw
                    # pooling layers: full-resolution signs go into the (unused) int16 map's memory
                    out = L.y.view(torch.int8) if l.pool else L.s8
                    self._call("zkb_bconv2d_fwd_i8_sign", zn.ptr(prev.s8), zn.ptr(L.w_fwd8), zn.ptr(L.thr), zn.ptr(out),
                               B, ih, iw, l.cin, l.cout, l.ksize)
                    if l.pool:
                        self._call("zkb_maxpool2_sign_i8", zn.ptr(out), B, oh, ow, L.cp, zn.ptr(L.s8))
                else:
                    self._call("zkb_bdense_fwd_i8_sign", zn.ptr(prev.s8), zn.ptr(L.w_fwd8), zn.ptr(L.thr), zn.ptr(L.s8),
                               L.cp, B, l.cin, L.cp)
            prev = L

    def last_loss(self) -> float:
        """Mean cross-entropy of the last step over this replica's batch (synchronises)."""
        return float(self.acc[4 * self.n_beta].item()) / self.B

    def last_accuracy(self) -> float:
        return float(self.correct.item()) / self.B

    def logits(self) -> np.ndarray:
        return self.L[-1].logits[:, : self.num_classes].detach().cpu().numpy()
