"""Keras-shaped model front end over :class:`zookeeper_b200.engine.BinaryNetEngine`.

``Sequential([...])`` accepts the layer list of the reference's ``BinaryNet.build``
(examples/larq_experiment.py:59-103), checks that it is a stack the CUDA engine implements and
lowers it to engine blocks; ``compile`` / ``summary`` / ``fit`` / ``evaluate`` keep the call shapes
``run()`` uses (:141-153), so the example's train loop reads like the reference's.
"""
from __future__ import annotations

import math
import os
import time
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import image as zimage
from . import layers as zl
from .data import DataPlan
from .optimizers import Adam


class Model:
    """Base type for the ``model`` field of an Experiment."""


def _lower_layers(layers: Sequence) -> Tuple[Tuple[int, int, int], list, int]:
    from ..engine import LayerSpec

    if not layers:
        raise ValueError("Sequential needs at least one layer")
    first = layers[0]
    if not isinstance(first, zl.QuantConv2D) or first.input_shape is None:
        raise NotImplementedError("the first layer must be a QuantConv2D with input_shape")
    input_shape = tuple(int(v) for v in first.input_shape)
    h, w, c = input_shape
    blocks: List[LayerSpec] = []
    i, flat, width = 0, False, c
    while i < len(layers):
        l = layers[i]
        if isinstance(l, zl.Flatten):
            flat, width = True, h * w * width
            i += 1
            continue
        if isinstance(l, zl.Activation):
            if l.activation != "softmax" or i != len(layers) - 1:
                raise NotImplementedError("only a terminal softmax activation is supported")
            i += 1
            continue
        if not isinstance(l, (zl.QuantConv2D, zl.QuantDense)):
            raise NotImplementedError(f"unexpected layer {l!r}: blocks are QuantConv2D|QuantDense [MaxPool2D] BatchNormalization")
        if l.kernel_quantizer != "ste_sign" or l.kernel_constraint != "weight_clip" or l.use_bias:
            raise NotImplementedError("layers must use kernel_quantizer='ste_sign', kernel_constraint='weight_clip', use_bias=False")
        quant_in = l.input_quantizer == "ste_sign"
        if l.input_quantizer not in (None, "ste_sign"):
            raise NotImplementedError(f"input_quantizer {l.input_quantizer!r} is not implemented")
        if quant_in == (len(blocks) == 0):
            raise NotImplementedError("exactly the first layer must have no input quantizer")
        if isinstance(l, zl.QuantConv2D):
            if flat:
                raise NotImplementedError("conv after Flatten")
            if l.kernel_size[0] != l.kernel_size[1] or l.strides != (1, 1):
                raise NotImplementedError("square kernels with stride 1 only")
            ks = l.kernel_size[0]
            pad = l.padding.lower()
            spec = LayerSpec("conv", width, l.filters, ks, pad, quant_in, False)
            if pad == "valid":
                h, w = h - ks + 1, w - ks + 1
            elif pad != "same":
                raise NotImplementedError(f"padding {l.padding!r}")
            width = l.filters
        else:
            if not flat:
                raise NotImplementedError("QuantDense needs a Flatten before it")
            spec = LayerSpec("dense", width, l.units, 1, "valid", quant_in, False)
            width = l.units
        i += 1
        if i < len(layers) and isinstance(layers[i], zl.MaxPool2D):
            mp = layers[i]
            if spec.kind != "conv" or mp.pool_size != (2, 2) or mp.strides != (2, 2):
                raise NotImplementedError("only MaxPool2D((2, 2), strides=(2, 2)) after a conv is supported")
            spec.pool = True
            h, w = h // 2, w // 2
            i += 1
        if i >= len(layers) or not isinstance(layers[i], zl.BatchNormalization):
            raise NotImplementedError("every conv/dense block must end in BatchNormalization(scale=False)")
        bn = layers[i]
        if bn.scale or not bn.center or bn.momentum != 0.99 or bn.epsilon != 1e-3:
            raise NotImplementedError("BatchNormalization(scale=False) with keras default momentum/epsilon only")
        i += 1
        blocks.append(spec)
    if blocks[-1].kind != "dense":
        raise NotImplementedError("the stack must end in a QuantDense classifier")
    return input_shape, blocks, blocks[-1].cout


def _lower_plan(plan: DataPlan, input_shape) -> Tuple[zimage.FusedPreprocess, bool]:
    """Run the mapped Preprocessing once on symbolic features and lower the recorded ops."""
    if plan.map_fn is None:
        raise ValueError("the dataset plan has no .map(preprocessing)")
    src = tuple(int(v) for v in plan.images.shape[1:])
    features = {"image": zimage.ImageSpec(src), "label": "label"}
    inp, out = plan.map_fn(features)
    if not isinstance(inp, zimage.ImageSpec):
        raise TypeError("Preprocessing.input must return the symbolic image it was given, transformed")
    if out != "label":
        raise NotImplementedError("Preprocessing.output must return data['label'] (sparse labels)")
    if tuple(inp.shape) != tuple(input_shape):
        raise ValueError(f"preprocessing produces shape {inp.shape}, the model expects {tuple(input_shape)}")
    return zimage.lower(inp, src[:2]), plan.training


class Sequential(Model):
    def __init__(self, layers: Sequence):
        self.layers = list(layers)
        self.input_shape, self.blocks, self.num_classes = _lower_layers(self.layers)
        self.optimizer: Optional[Adam] = None
        self.loss = None
        self.metrics: List[str] = []
        self.engine = None              # the engine that holds the model's state (first one built)
        self._engines = {}              # per-rank batch size -> engine; all alias self.engine's state
        self.history: Dict[str, List[float]] = {}

    # ------------------------------------------------------------------ keras-shaped API
    def compile(self, optimizer="adam", loss=None, metrics=None):
        if isinstance(optimizer, str):
            if optimizer.lower() != "adam":
                raise NotImplementedError(f"optimizer {optimizer!r}: only Adam is implemented")
            optimizer = Adam()
        if not isinstance(optimizer, Adam):
            raise NotImplementedError("only zookeeper_b200.tf.optimizers.Adam is implemented")
        if loss != "sparse_categorical_crossentropy":
            raise NotImplementedError("loss must be 'sparse_categorical_crossentropy'")
        for m in metrics or []:
            if m != "accuracy":
                raise NotImplementedError(f"metric {m!r}: only 'accuracy' is implemented")
        self.optimizer, self.loss, self.metrics = optimizer, loss, list(metrics or [])

    def count_params(self) -> int:
        n = 0
        for b in self.blocks:
            n += (b.ksize * b.ksize if b.kind == "conv" else 1) * b.cin * b.cout + b.cout
        return n

    def summary(self, print_fn=print) -> None:
        """Per-block table in the spirit of ``lq.models.summary`` (examples/larq_experiment.py:145)."""
        h, w, _ = self.input_shape
        rows, total_macs = [], 0
        for i, b in enumerate(self.blocks):
            if b.kind == "conv":
                if b.padding == "valid":
                    h, w = h - b.ksize + 1, w - b.ksize + 1
                macs = h * w * b.ksize * b.ksize * b.cin * b.cout
                shape = (h // 2, w // 2, b.cout) if b.pool else (h, w, b.cout)
                if b.pool:
                    h, w = h // 2, w // 2
            else:
                macs, shape = b.cin * b.cout, (b.cout,)
            total_macs += macs
            kind = ("quant_conv2d" if b.kind == "conv" else "quant_dense") + ("+pool" if b.pool else "") + "+bn"
            rows.append((f"{kind}_{i}", "1-bit" if b.quant_input else "-", str(shape),
                         (b.ksize * b.ksize if b.kind == "conv" else 1) * b.cin * b.cout + b.cout, macs))
        print_fn(f"{'Layer':24s}{'Input prec.':>12s}{'Outputs':>18s}{'# params':>12s}{'MACs':>14s}")
        for r in rows:
            print_fn(f"{r[0]:24s}{r[1]:>12s}{r[2]:>18s}{r[3]:>12d}{r[4]:>14d}")
        print_fn(f"Total params {self.count_params()}  |  MACs/image {total_macs}")

    # ------------------------------------------------------------------ engine plumbing
    def _dist(self):
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
        return 0, 1

    def _ensure_engine(self, global_batch: int):
        """The engine for this (global) batch size.  Activation buffers are sized per batch, so
        each batch size gets its own engine -- but all of them ALIAS one set of parameter,
        Adam-moment and moving-statistics arenas and one step counter (``share_state_with``):
        validating at another batch size than training sees the weights being trained, not a copy."""
        from ..engine import BinaryNetEngine

        rank, world = self._dist()
        if global_batch % world:
            raise ValueError(f"batch size {global_batch} is not divisible by the {world} replicas")
        per_rank = global_batch // world
        if per_rank in self._engines:
            return self._engines[per_rank]
        if self.optimizer is None:
            raise RuntimeError("call compile() before fit()/evaluate()")
        opt = self.optimizer
        eng = BinaryNetEngine(self.blocks, self.input_shape, per_rank, self.num_classes,
                              learning_rate=opt.learning_rate, beta_1=opt.beta_1, beta_2=opt.beta_2,
                              epsilon=opt.epsilon, share_state_with=self.engine)
        if self.engine is None:
            eng.init_glorot(0)
            self.engine = eng
        self._engines[per_rank] = eng
        return eng

    def _sum_over_replicas(self, *tensors):
        """Loss / accuracy counters are sums over this replica's samples: add the other replicas'."""
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            for t in tensors:
                dist.all_reduce(t)

    @staticmethod
    def _resident(plan: DataPlan, device):
        import torch

        imgs = plan.images if isinstance(plan.images, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(plan.images))
        labs = plan.labels if isinstance(plan.labels, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(plan.labels))
        return imgs.to(device, non_blocking=True), labs.to(device=device, dtype=torch.int64, non_blocking=True)

    def fit(self, train_data: DataPlan, epochs: int = 1, steps_per_epoch: Optional[int] = None,
            validation_data: Optional[DataPlan] = None, validation_steps: Optional[int] = None, verbose: int = 1):
        import torch

        spec, training = _lower_plan(train_data, self.input_shape)
        if not training:
            raise ValueError("map the training data with partial(preprocessing, training=True)")
        if validation_data is not None:
            _lower_plan(validation_data, self.input_shape)
        if self.optimizer is None:
            raise RuntimeError("call compile() before fit()")
        if epochs <= 0:
            # like keras: nothing to run (the reference's CI smoke, unittest.yml:41-42); everything
            # up to here -- configuration, model lowering, preprocessing lowering -- has been checked
            self.history = {"loss": [], "accuracy": [], "images_per_sec": []}
            return self.history
        eng = self._ensure_engine(train_data.batch_size)
        rank, world = self._dist()
        d_img, d_lab = self._resident(train_data, eng.device)
        if steps_per_epoch is None:
            steps_per_epoch = len(train_data) // train_data.batch_size
        stream = train_data.index_stream()
        nb4 = 4 * eng.n_beta
        pad = spec.pad_size if spec.random_crop else 0
        seed = train_data.seed + rank
        # CUDA-graph replay of the step on one GPU only: a captured step that holds NCCL work kept
        # the process-group teardown from returning at the end of an 8-GPU run (r2), so data-parallel
        # runs stay on per-kernel launches (at their per-GPU batch sizes the step is not launch-bound)
        use_graph = os.environ.get("ZKB_NO_GRAPH", "0") != "1" and world == 1
        captured = False
        hist = {"loss": [], "accuracy": [], "images_per_sec": []}
        if validation_data is not None:
            hist.update({"val_loss": [], "val_accuracy": []})
        for epoch in range(epochs):
            run_loss = torch.zeros(1, dtype=torch.float64, device=eng.device)
            run_ok = torch.zeros(1, dtype=torch.int64, device=eng.device)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            # one upload of the epoch's sample indices; a step then takes its row on the device
            epoch_idx = np.stack([next(stream)[rank::world] for _ in range(steps_per_epoch)]).astype(np.int32)
            d_epoch_idx = torch.as_tensor(epoch_idx).to(eng.device, non_blocking=True)
            for k in range(steps_per_epoch):
                if captured:
                    eng.train_step_graph(d_epoch_idx[k])
                else:
                    eng.train_step(d_img, d_lab, d_epoch_idx[k], pad_size=pad, seed=seed)
                    if use_graph:
                        # after one ordinary step (first-use set-up done) replay a captured graph:
                        # a step becomes one launch instead of ~90
                        try:
                            eng.capture_train_step(d_img, d_lab, pad, seed)
                            captured = True
                        except Exception as exc:  # keep training without the graph
                            use_graph = False
                            if rank == 0:
                                print(f"CUDA-graph capture unavailable ({exc}); continuing with per-kernel launches")
                run_loss += eng.acc[nb4:nb4 + 1]
                run_ok += eng.correct
            self._sum_over_replicas(run_loss, run_ok)     # the logged metrics cover the GLOBAL batches
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            seen = steps_per_epoch * eng.B * world
            hist["loss"].append(float(run_loss.item()) / max(seen, 1))
            hist["accuracy"].append(float(run_ok.item()) / max(seen, 1))
            hist["images_per_sec"].append(seen / dt if dt > 0 else float("nan"))
            msg = (f"Epoch {epoch + 1}/{epochs} - {dt:.2f}s - loss: {hist['loss'][-1]:.4f} - "
                   f"accuracy: {hist['accuracy'][-1]:.4f} - {hist['images_per_sec'][-1]:.0f} images/s")
            if validation_data is not None:
                vl, va = self.evaluate(validation_data, validation_steps, verbose=0)
                hist["val_loss"].append(vl)
                hist["val_accuracy"].append(va)
                msg += f" - val_loss: {vl:.4f} - val_accuracy: {va:.4f}"
            if verbose and rank == 0:
                print(msg, flush=True)
        self.history = hist
        return hist

    def save_weights(self, path: str) -> None:
        """Checkpoint (weights in Keras HWIO / [K, N] layout, moving stats, Adam state) as .npz."""
        if self.engine is None:
            raise RuntimeError("nothing to save: the model has not been fitted or evaluated yet")
        self.engine.average_moving_stats()     # per-replica BatchNorm: one checkpoint whatever the rank
        self.engine.save_checkpoint(path)

    def load_weights(self, path: str, batch_size: int) -> None:
        self._ensure_engine(batch_size).load_checkpoint(path)

    def evaluate(self, data: DataPlan, steps: Optional[int] = None, verbose: int = 1):
        import torch

        eng = self._ensure_engine(data.batch_size)
        rank, world = self._dist()
        _lower_plan(data, self.input_shape)
        eng.average_moving_stats()             # per-replica BatchNorm: validate with the replicas' mean
        d_img, d_lab = self._resident(data, eng.device)
        if steps is None:
            steps = len(data) // data.batch_size
        stream = data.index_stream()
        nb4 = 4 * eng.n_beta
        run_loss = torch.zeros(1, dtype=torch.float64, device=eng.device)
        run_ok = torch.zeros(1, dtype=torch.int64, device=eng.device)
        for _ in range(steps):
            idx = next(stream)[rank::world]
            d_idx = torch.as_tensor(idx.astype(np.int32)).to(eng.device, non_blocking=True)
            eng.eval_step(d_img, d_lab, d_idx)
            run_loss += eng.acc[nb4:nb4 + 1]
            run_ok += eng.correct
        self._sum_over_replicas(run_loss, run_ok)
        seen = max(steps * eng.B * world, 1)
        loss, acc = float(run_loss.item()) / seen, float(run_ok.item()) / seen
        if verbose and rank == 0:
            print(f"evaluate - loss: {loss:.4f} - accuracy: {acc:.4f}", flush=True)
        return loss, acc
