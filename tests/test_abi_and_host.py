"""CPU-side checks: the C-ABI library builds, loads and exports every symbol include/zkb200.h
declares (no compute calls without a GPU); the host logic of the tf-shaped front end (layer lowering,
preprocessing lowering, dataset plans, example CLI); product code never imports the oracle."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "zkb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(zkb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(native_lib):
    from zookeeper_b200 import native

    declared = _declared_symbols()
    assert len(declared) >= 25
    missing = [s for s in declared if not hasattr(native_lib, s)]
    assert not missing, f"declared in zkb200.h but not exported: {missing}"
    # and the ctypes table covers every declared compute entry point
    untyped = [s for s in declared if s not in native.SIGNATURES
               and s not in ("zkb_abi_version", "zkb_last_error_string", "zkb_bn_train_sign_static_max_channels",
                            "zkb_bconv2d_fwd_i8_pool_ok")]
    assert not untyped, f"no ctypes signature for: {untyped}"
    assert native_lib.zkb_abi_version() == 2
    # pure host query: which pooled maps the fused conv+pool forward tiles (even sizes only)
    ok = native_lib.zkb_bconv2d_fwd_i8_pool_ok
    assert [ok(26, 26, 128), ok(6, 6, 512), ok(30, 30, 128), ok(13, 13, 256), ok(7, 7, 512), ok(26, 26, 96)] == [1, 1, 1, 0, 0, 0]


def test_argument_errors_are_reported_before_any_launch(native_lib):
    from zookeeper_b200 import native as zn

    with pytest.raises(zn.ZkbError, match="null pointer") as e:
        zn.call("zkb_preprocess_u8", 0, 0, 0, 0, 0, 0, 0, 4, 28, 28, 1, 28, 28, 32, 1, 0, 10, 0, 0, 0, 0)
    assert e.value.status < 0
    with pytest.raises(zn.ZkbError, match="multiple of 32"):
        zn.call("zkb_sign_pack", 16, 16, 33, 0, 0)
    with pytest.raises(zn.ZkbError, match="symbol is not exported"):
        zn.call("zkb_does_not_exist")


def test_engine_refuses_to_run_without_cuda():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        BinaryNetEngine(binarynet_layers(), (28, 28, 1), 8, 10)
    from zookeeper_b200.bireal import BiRealUnitsEngine

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        BiRealUnitsEngine(("plain", "down"), 128, (14, 14), 4, 10)
    from zookeeper_b200.alexnet import BinaryAlexNetEngine
    from zookeeper_b200.bireal import BiRealNetEngine, bireal18_plan

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        BiRealNetEngine(bireal18_plan(), (224, 224), 4, 1000)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        BinaryAlexNetEngine((224, 224), 4, 1000)
    assert bireal18_plan().count("down") == 3 and len(bireal18_plan()) == 16


def test_product_code_never_imports_the_oracle():
    offenders = []
    for base, _, files in os.walk(os.path.join(ROOT, "zookeeper_b200")):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(base, f)).read()
                if re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M):
                    offenders.append(f)
    assert not offenders


# ------------------------------------------------------------------------------------ front end
def test_sequential_lowering_matches_reference_topology():
    from zookeeper_b200.engine import binarynet_layers
    from zookeeper_b200.tf import layers as L
    from zookeeper_b200.tf.models import Sequential

    kw = dict(input_quantizer="ste_sign", kernel_quantizer="ste_sign", kernel_constraint="weight_clip", use_bias=False)
    stack = [
        L.QuantConv2D(128, (3, 3), kernel_quantizer="ste_sign", kernel_constraint="weight_clip", use_bias=False,
                      input_shape=(32, 32, 3)),
        L.BatchNormalization(scale=False),
        L.QuantConv2D(128, 3, padding="same", **kw), L.MaxPool2D((2, 2), (2, 2)), L.BatchNormalization(scale=False),
        L.QuantConv2D(256, 3, padding="same", **kw), L.BatchNormalization(scale=False),
        L.QuantConv2D(256, 3, padding="same", **kw), L.MaxPool2D((2, 2), (2, 2)), L.BatchNormalization(scale=False),
        L.QuantConv2D(512, 3, padding="same", **kw), L.BatchNormalization(scale=False),
        L.QuantConv2D(512, 3, padding="same", **kw), L.MaxPool2D((2, 2), (2, 2)), L.BatchNormalization(scale=False),
        L.Flatten(),
        L.QuantDense(1024, **kw), L.BatchNormalization(scale=False),
        L.QuantDense(1024, **kw), L.BatchNormalization(scale=False),
        L.QuantDense(10, **kw), L.BatchNormalization(scale=False),
        L.Activation("softmax"),
    ]
    m = Sequential(stack)
    assert m.blocks == binarynet_layers((32, 32, 3), 128, 1024, 10)     # CIFAR-shaped config 3
    assert m.count_params() == 10352000 + 3850                          # kernels + betas (appendix B)
    lines = []
    m.summary(print_fn=lines.append)
    assert "MACs/image 514072064" in lines[-1]

    with pytest.raises(NotImplementedError, match="BatchNormalization"):
        Sequential(stack[:1] + stack[2:])
    bad = list(stack)
    bad[2] = L.QuantConv2D(128, 3, padding="same", kernel_quantizer="ste_sign", kernel_constraint="weight_clip", use_bias=True)
    with pytest.raises(NotImplementedError, match="use_bias=False"):
        Sequential(bad)
    m.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    with pytest.raises(NotImplementedError):
        m.compile(optimizer="sgd", loss="sparse_categorical_crossentropy")


def test_preprocessing_lowering_and_training_kwarg():
    from functools import partial

    from zookeeper_b200.core import Field, component, configure
    from zookeeper_b200.tf import Preprocessing
    from zookeeper_b200.tf import image as zi

    @component
    class PadCropAndFlip(Preprocessing):
        pad_size: int = Field(32)
        input_shape = Field((28, 28, 1))

        def input(self, data, training):
            image = data["image"]
            if training:
                image = zi.resize_with_crop_or_pad(image, self.pad_size, self.pad_size)
                image = zi.random_crop(image, self.input_shape)
                image = zi.random_flip_left_right(image)
            else:
                image = zi.resize_with_crop_or_pad(image, *self.input_shape[:2])
            return zi.cast(image, "float32") / (255.0 / 2.0) - 1.0

        def output(self, data):            # no `training` parameter: must not be passed one
            return data["label"]

    p = PadCropAndFlip()
    configure(p, {})
    feats = {"image": zi.ImageSpec((28, 28, 1)), "label": "label"}
    inp, out = partial(p, training=True)(feats)
    spec = zi.lower(inp, (28, 28))
    assert (spec.pad_size, spec.random_crop, spec.random_flip, spec.out_hw) == (32, True, True, (28, 28))
    assert out == "label"
    inp, _ = p(feats)                      # training defaults to False (preprocessing.py:59)
    spec = zi.lower(inp, (28, 28))
    assert (spec.pad_size, spec.random_crop, spec.random_flip) == (0, False, False)
    with pytest.raises(NotImplementedError):
        zi.lower(zi.cast(feats["image"]) / 255.0, (28, 28))
    with pytest.raises(NotImplementedError, match="Must be implemented"):
        Preprocessing.input(p, feats, False) if False else Preprocessing().input(feats, False)


def test_dataset_contract_and_plan_chain():
    from zookeeper_b200.core import Field, component, configure
    from zookeeper_b200.tf import Dataset, SyntheticDataset
    from zookeeper_b200.tf.data import DataPlan

    @component
    class Tiny(SyntheticDataset):
        num_train_examples = Field(100)
        num_validation_examples = Field(0)

    d = Tiny()
    configure(d, {})
    plan, n = d.train()
    assert n == 100 and plan.images.shape == (100, 28, 28, 1) and plan.images.dtype == np.uint8
    with pytest.raises(ValueError, match="is not configured with validation data"):
        d.validation()
    with pytest.raises(ValueError, match="is not configured with test data"):
        d.test()
    with pytest.raises(TypeError):
        Dataset()                           # abstract

    chained = plan.cache().shuffle(40, seed=3).repeat().map(lambda x: x).batch(16)
    assert isinstance(chained, DataPlan) and chained.batch_size == 16 and chained.repeated
    it = chained.index_stream()
    first_epoch = np.concatenate([next(it) for _ in range(6)])          # 96 of 100 samples
    assert len(set(first_epoch.tolist())) == 96 and first_epoch.max() < 100
    # streaming shuffle with buffer 40: element i cannot appear before position i - 40 + 1
    pos = {int(v): k for k, v in enumerate(first_epoch)}
    assert all(k >= v - 39 for v, k in pos.items())
    assert not np.array_equal(first_epoch, np.arange(96))
    # repeat() before batch(): the 4-sample tail of pass 1 opens the next batch (tf.data semantics)
    seventh = next(it)
    assert set(seventh[:4].tolist()) == set(range(100)) - set(first_epoch.tolist())
    two_passes = np.concatenate([first_epoch, seventh] + [next(it) for _ in range(5)] + [next(it)[:8]])
    assert len(two_passes) == 200 and sorted(two_passes.tolist()) == sorted(list(range(100)) * 2)
    seq = plan.batch(32)
    batches = list(seq.index_stream())
    assert len(batches) == 3 and np.array_equal(batches[0], np.arange(32))   # remainder dropped


def test_example_cli_configures_and_lowers_without_a_gpu():
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "larq_experiment.py"), "binary_net_mnist",
                        "epochs=0", "dataset.num_train_examples=256", "dataset.num_validation_examples=128"],
                       capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Total params 10353546" in r.stdout and "MACs/image 383158784" in r.stdout
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "larq_experiment.py"), "BinaryNetMnist",
                        "epochs=0", "no_such_field=1"], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode != 0 and "does not correspond to any field" in (r.stderr + r.stdout + str(r))


def test_bench_reference_arm_prints_one_contract_line():
    """`bench.py --impl reference` (the CPU restatement timed on host cores) prints exactly one JSON
    line with the keys the driver reads; no GPU needed."""
    import json

    env = dict(os.environ, PYTHONPATH=ROOT, ZKB_REF_SKIP_B128="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "1", "--batch", "32"], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "BinaryNet train images/sec" and d["unit"] == "images/s"
    assert d["higher_is_better"] is True and d["steps"] == 1 and d["warmup"] == 1 and d["value"] > 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["e2e"]["value"] == d["value"] and "workload" in d["config"]
    # the arm runs the batch it was asked for (the GPU arm's), and says so
    assert d["config"]["global_batch"] == 32 and "batch 32/GPU" in d["config"]["workload"]
    assert "batch 32" in d["cpu_baseline"]["sample"] and d["reference_run"]["batch"] == 32
    # the same config object the GPU arm prints for this command line
    assert set(d["config"]) == {"workload", "global_batch", "parallelism", "batchnorm", "launch", "l2"}
    assert d["config"]["parallelism"] == "dp1" and d["config"]["batchnorm"] == "per-replica"
    # BASELINE.json configs[2] (CIFAR-10-shaped), strong split of its batch over 8 ranks: 256 / 8 per rank
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--config", "cifar", "--scaling", "strong", "--gpus", "8"], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads([ln for ln in r.stdout.splitlines() if ln.strip()][0])
    assert "CIFAR-10-shaped 32x32x3" in d["config"]["workload"] and d["scaling"] == "strong"
    assert d["config"]["global_batch"] == 256 and d["config"]["parallelism"] == "dp8" and d["reference_run"]["batch"] == 32
