"""Keras-shaped handles on the two larq-zoo models the engines implement (SURVEY.md section 8f.2;
BASELINE.json configs[3], [4]): ``BinaryAlexNet(...)`` and ``BiRealNet18(...)`` return a model with the
``compile`` / ``summary`` / ``fit`` call shapes the reference's ``run()`` uses
(/root/reference/examples/larq_experiment.py:141-153), over ``zookeeper_b200.alexnet.BinaryAlexNetEngine``
and ``zookeeper_b200.bireal.BiRealNetEngine``.  The reference reaches such models through a
``@factory`` that builds them from larq / larq-zoo (examples/larq_experiment.py:41-57 is the hook);
``examples/larq_zoo_experiment.py`` is that hook here.

Training only: the zoo engines have no inference path yet, so ``fit`` refuses ``validation_data`` and
there is no ``evaluate``.  Everything up to the first train step -- configuration, model description,
parameter counts, the lowering of the mapped ``Preprocessing`` to the fused kernel -- runs without a
GPU (``epochs=0`` is the reference CI's smoke, .github/workflows/unittest.yml:41-42).
"""
from __future__ import annotations

import time
from typing import Dict, List, Optional, Tuple

import numpy as np

from .data import DataPlan
from .models import Model, _lower_plan
from .optimizers import Adam

ALEXNET_FEATURES = (64, 192, 384, 384, 256)
ALEXNET_KSIZES = (11, 5, 3, 3, 3)
ALEXNET_STRIDES = (4, 1, 1, 1, 1)
ALEXNET_POOLS = (True, True, False, False, True)


def bireal18_plan() -> List[str]:
    return ["plain"] * 4 + (["down"] + ["plain"] * 3) * 3


def _alexnet_rows(hw, num_classes, dense_units):
    """(name, input precision, output shape, params, MACs) per block; the layer shapes of
    zookeeper_b200/alexnet.py."""
    rows, (h, w), c = [], hw, 3
    for i, (f, k, s, pool) in enumerate(zip(ALEXNET_FEATURES, ALEXNET_KSIZES, ALEXNET_STRIDES, ALEXNET_POOLS)):
        ch, cw = -(-h // s), -(-w // s)
        oh, ow = ((ch - 3) // 2 + 1, (cw - 3) // 2 + 1) if pool else (ch, cw)
        if ch < 3 or cw < 3:
            raise ValueError(f"image {hw[0]}x{hw[1]} is too small for the three 3x3/2 pools")
        rows.append((f"quant_conv2d{'+pool' if pool else ''}+bn_{i}", "-" if i == 0 else "1-bit", (oh, ow, f),
                     k * k * c * f + f, ch * cw * k * k * c * f))
        h, w, c = oh, ow, f
    flat = h * w * c
    for j, (kin, kout) in enumerate(((flat, dense_units), (dense_units, dense_units), (dense_units, num_classes))):
        rows.append((f"quant_dense+bn_{5 + j}", "1-bit", (kout,), kin * kout + kout, kin * kout))
    return rows


def _bireal_rows(hw, num_classes, filters=64):
    h1, w1 = -(-hw[0] // 2), -(-hw[1] // 2)
    h, w, c = (h1 + 1) // 2, (w1 + 1) // 2, filters
    rows = [("conv2d_7x7/2+bn+pool_stem", "-", (h, w, c), 49 * 3 * c + 2 * c, h1 * w1 * 147 * c)]
    for u, kind in enumerate(bireal18_plan()):
        if kind == "down":
            if h % 2 or w % 2:
                raise ValueError(f"a down-sampling unit needs an even map, got {h}x{w}")
            h, w, co = h // 2, w // 2, 2 * c
            params = 9 * c * co + 2 * co + c * co + 2 * co          # + 1x1 float shortcut conv and its BatchNorm
            macs = h * w * (9 * c * co + c * co)
        else:
            co = c
            params, macs = 9 * c * co + 2 * co, h * w * 9 * c * co
        rows.append((f"residual_unit_{kind}_{u}", "1-bit", (h, w, co), params, macs))
        c = co
    rows.append(("global_avgpool+dense_head", "-", (num_classes,), c * num_classes + num_classes, c * num_classes))
    return rows


class ZooModel(Model):
    """``kind``: "alexnet" or "bireal18".  ``input_shape``: (H, W, 3)."""

    def __init__(self, kind: str, input_shape: Tuple[int, int, int], num_classes: int, dense_units: int = 4096):
        if kind not in ("alexnet", "bireal18"):
            raise ValueError('kind must be "alexnet" or "bireal18"')
        if len(input_shape) != 3 or int(input_shape[2]) != 3:
            raise NotImplementedError("the zoo models take 3-channel images")
        self.kind, self.input_shape = kind, tuple(int(v) for v in input_shape)
        self.num_classes, self.dense_units = int(num_classes), int(dense_units)
        hw = self.input_shape[:2]
        self.rows = _alexnet_rows(hw, self.num_classes, self.dense_units) if kind == "alexnet" else _bireal_rows(hw, self.num_classes)
        self.optimizer: Optional[Adam] = None
        self.loss = None
        self.metrics: List[str] = []
        self.engine = None
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
        return sum(r[3] for r in self.rows)

    def summary(self, print_fn=print) -> None:
        """Per-block table in the spirit of ``lq.models.summary`` (examples/larq_experiment.py:145)."""
        print_fn(f"{'Layer':34s}{'Input prec.':>12s}{'Outputs':>18s}{'# params':>12s}{'MACs':>14s}")
        for r in self.rows:
            print_fn(f"{r[0]:34s}{r[1]:>12s}{str(r[2]):>18s}{r[3]:>12d}{r[4]:>14d}")
        print_fn(f"Total params {self.count_params()}  |  MACs/image {sum(r[4] for r in self.rows)}")

    # ------------------------------------------------------------------ engine plumbing
    @staticmethod
    def _dist():
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
        return 0, 1

    def _glorot_params(self, eng, seed: int = 0) -> Dict[str, np.ndarray]:
        """Glorot-uniform kernels, BatchNorm at identity -- the same on every replica."""
        rng = np.random.default_rng(seed)
        p: Dict[str, np.ndarray] = {}

        def glorot(shape, fan_in, fan_out):
            lim = np.sqrt(6.0 / (fan_in + fan_out))
            return rng.uniform(-lim, lim, shape).astype(np.float32)

        if self.kind == "alexnet":
            for L in eng.L:
                rf = L.k * L.k if L.kind == "conv" else 1
                p[f"kernel{L.idx}"] = glorot(eng._kshape(L), rf * L.cin, rf * L.cout)
                p[f"beta{L.idx}"] = np.zeros(L.cout, np.float32)
            return p
        c = eng.F
        p["stem_kernel"] = glorot((eng.ks, eng.ks, 3, c), eng.ks * eng.ks * 3, eng.ks * eng.ks * c)
        p["stem_gamma"], p["stem_beta"] = np.ones(c, np.float32), np.zeros(c, np.float32)
        for u, X in enumerate(eng.body.units):
            ci, co = X.in_shape[2], X.out_shape[2]
            p[f"kernel{u}"] = glorot((3, 3, ci, co), 9 * ci, 9 * co)
            p[f"gamma{u}"], p[f"beta{u}"] = np.ones(co, np.float32), np.zeros(co, np.float32)
            if X.kind == "down":
                p[f"sc_kernel{u}"] = glorot((ci, co), ci, co)
                p[f"sc_gamma{u}"], p[f"sc_beta{u}"] = np.ones(co, np.float32), np.zeros(co, np.float32)
        oc = eng.body.out_c
        p["dense_kernel"] = glorot((oc, self.num_classes), oc, self.num_classes)
        p["dense_bias"] = np.zeros(self.num_classes, np.float32)
        return p

    def _ensure_engine(self, global_batch: int):
        rank, world = self._dist()
        if global_batch % world:
            raise ValueError(f"batch size {global_batch} is not divisible by the {world} replicas")
        per_rank = global_batch // world
        if self.engine is not None:
            if self.engine.B != per_rank:
                raise NotImplementedError("the zoo models keep one engine: the batch size cannot change between fits")
            return self.engine
        if self.optimizer is None:
            raise RuntimeError("call compile() before fit()")
        opt = self.optimizer
        kw = dict(learning_rate=opt.learning_rate, beta_1=opt.beta_1, beta_2=opt.beta_2, epsilon=opt.epsilon)
        hw = self.input_shape[:2]
        if self.kind == "alexnet":
            from ..alexnet import BinaryAlexNetEngine

            eng = BinaryAlexNetEngine(hw, per_rank, self.num_classes, dense_units=self.dense_units, **kw)
        else:
            from ..bireal import BiRealNetEngine

            eng = BiRealNetEngine(bireal18_plan(), hw, per_rank, self.num_classes, **kw)
        eng.load_params(self._glorot_params(eng))
        self.engine = eng
        return eng

    def fit(self, train_data: DataPlan, epochs: int = 1, steps_per_epoch: Optional[int] = None,
            validation_data: Optional[DataPlan] = None, validation_steps: Optional[int] = None, verbose: int = 1):
        spec, training = _lower_plan(train_data, self.input_shape)
        if not training:
            raise ValueError("map the training data with partial(preprocessing, training=True)")
        if validation_data is not None:
            raise NotImplementedError("the zoo models have no inference path yet: fit() without validation_data")
        if self.optimizer is None:
            raise RuntimeError("call compile() before fit()")
        if epochs <= 0:      # like keras: nothing to run; everything up to here has been checked
            self.history = {"loss": [], "accuracy": [], "images_per_sec": []}
            return self.history
        import torch

        from .. import native as zn

        eng = self._ensure_engine(train_data.batch_size)
        rank, world = self._dist()
        dev = eng.device
        d_img = torch.as_tensor(np.ascontiguousarray(train_data.images)).to(dev)
        d_lab = torch.as_tensor(np.ascontiguousarray(train_data.labels)).to(device=dev, dtype=torch.int64)
        _, IH, IW, C = d_img.shape
        H, W, _ = self.input_shape
        pad = spec.pad_size if spec.random_crop else 0
        seed = train_data.seed + rank
        if steps_per_epoch is None:
            steps_per_epoch = len(train_data) // train_data.batch_size
        stream_idx = train_data.index_stream()
        # where the engine keeps the step's loss sum and hit count
        host = eng if self.kind == "alexnet" else eng.body
        lab_out = torch.empty(eng.B, dtype=torch.int32, device=dev)
        hist = {"loss": [], "accuracy": [], "images_per_sec": []}
        step = 0
        for epoch in range(epochs):
            run_loss = torch.zeros(1, dtype=torch.float64, device=dev)
            run_ok = torch.zeros(1, dtype=torch.int64, device=dev)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            epoch_idx = np.stack([next(stream_idx)[rank::world] for _ in range(steps_per_epoch)]).astype(np.int32)
            d_epoch_idx = torch.as_tensor(epoch_idx).to(dev)
            cuda_stream = torch.cuda.current_stream(dev).cuda_stream
            for k in range(steps_per_epoch):
                # the fused batch map of Preprocessing.__call__ (zookeeper/tf/preprocessing.py:59-75), into the engine's input buffer
                zn.call("zkb_preprocess_u8", zn.ptr(d_img), zn.ptr(d_lab), zn.ptr(d_epoch_idx[k]), 0, zn.ptr(eng.x), zn.ptr(lab_out),
                        0, eng.B, IH, IW, C, H, W, pad, 1, int(spec.binarize), self.num_classes, seed, step, 0, cuda_stream)
                eng.train_step(eng.x, lab_out)
                run_loss += host.acc[host.loss_off:host.loss_off + 1]
                run_ok += host.correct
                step += 1
            if world > 1:
                import torch.distributed as dist

                dist.all_reduce(run_loss)
                dist.all_reduce(run_ok)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            seen = steps_per_epoch * eng.B * world
            hist["loss"].append(float(run_loss.item()) / max(seen, 1))
            hist["accuracy"].append(float(run_ok.item()) / max(seen, 1))
            hist["images_per_sec"].append(seen / dt if dt > 0 else float("nan"))
            if verbose and rank == 0:
                print(f"Epoch {epoch + 1}/{epochs} - {dt:.2f}s - loss: {hist['loss'][-1]:.4f} - "
                      f"accuracy: {hist['accuracy'][-1]:.4f} - {hist['images_per_sec'][-1]:.0f} images/s", flush=True)
        self.history = hist
        return hist


def BinaryAlexNet(input_shape=(224, 224, 3), num_classes: int = 1000, dense_units: int = 4096) -> ZooModel:
    """larq_zoo.literature.BinaryAlexNet ([EXT], restated in oracle/alexnet_oracle.py)."""
    return ZooModel("alexnet", input_shape, num_classes, dense_units)


def BiRealNet18(input_shape=(224, 224, 3), num_classes: int = 1000) -> ZooModel:
    """larq_zoo.literature.BiRealNet ([EXT], restated in oracle/bireal_oracle.py)."""
    return ZooModel("bireal18", input_shape, num_classes)
