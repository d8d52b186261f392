"""The zookeeper-facing side of the two larq-zoo models (examples/larq_zoo_experiment.py,
zookeeper_b200/tf/zoo.py): what runs before the first train step needs no GPU -- configuration through
the unchanged CLI, the model description (parameter counts: BinaryAlexNet 61 838 248 trainable, the
figure larq-zoo documents as 61.8 M; Bi-Real-Net-18 11 689 512, ResNet-18's count), the lowering of the
mapped Preprocessing to the fused kernel, and the refusals (no inference path yet)."""
import os
import subprocess
import sys
from functools import partial

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("task,params,macs", [("BinaryAlexNetImagenet", 61838248, 841367552),
                                              ("bi_real_net18_imagenet", 11689512, 1814073344)])
def test_zoo_example_cli_configures_and_describes_without_a_gpu(task, params, macs):
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "larq_zoo_experiment.py"), task, "epochs=0",
                        "dataset.num_train_examples=8"], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert f"Total params {params}" in r.stdout and f"MACs/image {macs}" in r.stdout


def _plan(n=8, hw=(256, 256), training=True, out=(224, 224, 3)):
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import larq_zoo_experiment as ex
    from zookeeper_b200 import configure
    from zookeeper_b200.tf.data import DataPlan

    pre = ex.RandomCropAndFlip()
    configure(pre, {"input_shape": out})
    rng = np.random.default_rng(0)
    plan = DataPlan(rng.integers(0, 256, (n,) + hw + (3,), dtype=np.uint8), rng.integers(0, 10, (n,)).astype(np.int64), 10)
    fn = partial(pre, training=True) if training else pre
    return plan.cache().repeat().map(fn).batch(4)


def test_zoo_model_contract_and_refusals():
    from zookeeper_b200.tf import zoo
    from zookeeper_b200.tf.optimizers import Adam

    m = zoo.BinaryAlexNet((224, 224, 3), 1000)
    assert m.count_params() == 61838248
    assert zoo.BiRealNet18((224, 224, 3), 1000).count_params() == 11689512
    assert zoo.bireal18_plan().count("down") == 3
    with pytest.raises(RuntimeError, match="compile"):
        m.fit(_plan(), epochs=1)
    with pytest.raises(NotImplementedError, match="sparse_categorical_crossentropy"):
        m.compile(Adam(1e-3), loss="mse")
    with pytest.raises(NotImplementedError, match="only Adam"):
        m.compile("sgd", loss="sparse_categorical_crossentropy")
    m.compile(Adam(1e-3), loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    assert m.fit(_plan(), epochs=0) == {"loss": [], "accuracy": [], "images_per_sec": []}
    with pytest.raises(NotImplementedError, match="no inference path"):
        m.fit(_plan(), epochs=1, validation_data=_plan(training=False))
    with pytest.raises(ValueError, match="training=True"):
        m.fit(_plan(training=False), epochs=1)
    with pytest.raises(ValueError, match="the model expects"):
        m.fit(_plan(out=(192, 192, 3)), epochs=0)
    with pytest.raises(NotImplementedError, match="3-channel"):
        zoo.BinaryAlexNet((224, 224, 1), 10)
    with pytest.raises(ValueError, match="too small"):
        zoo.BinaryAlexNet((32, 32, 3), 10)
    lines = []
    zoo.BiRealNet18((64, 64, 3), 10).summary(lines.append)
    assert len(lines) == 2 + 18 and "residual_unit_down_4" in "\n".join(lines)


def test_zoo_fit_loop_on_a_stub_engine(monkeypatch):
    """The host loop of ``ZooModel.fit`` -- index stream, one fused-preprocessing call and one engine step
    per batch, loss / accuracy bookkeeping, history -- driven on the CPU with a stub in place of the CUDA
    engine and of the C-ABI call (the real pair is what scripts/bench_zoo.py runs on the GPU)."""
    import torch

    from zookeeper_b200 import native as zn
    from zookeeper_b200.tf import zoo
    from zookeeper_b200.tf.optimizers import Adam

    calls = []

    class StubEngine:
        B, device, loss_off = 4, torch.device("cpu"), 3

        def __init__(self):
            self.x = torch.zeros(4, 96, 96, 3)
            self.acc = torch.zeros(5, dtype=torch.float64)
            self.correct = torch.zeros(1, dtype=torch.int32)
            self.steps = 0

        def train_step(self, x, labels):
            assert x is self.x and labels.dtype == torch.int32 and labels.numel() == self.B
            self.steps += 1
            self.acc[self.loss_off] = 2.0 * self.B          # sum of the batch's losses
            self.correct[0] = 1

    stub = StubEngine()
    monkeypatch.setattr(zoo.ZooModel, "_ensure_engine", lambda self, gb: stub)
    monkeypatch.setattr(zn, "call", lambda name, *a: calls.append((name, a)))
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a, **k: type("S", (), {"cuda_stream": 0})())
    m = zoo.BinaryAlexNet((96, 96, 3), 10, dense_units=256)
    m.compile(Adam(1e-3), loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    hist = m.fit(_plan(n=8, hw=(128, 128), out=(96, 96, 3)), epochs=2, steps_per_epoch=3, verbose=0)
    assert stub.steps == 6 and len(calls) == 6 and all(c[0] == "zkb_preprocess_u8" for c in calls)
    a = calls[-1][1]
    # (pool, labels, index, aug, out, labels_out, onehot, B, IH, IW, C, H, W, pad, training, binarize, classes, seed, step, step_dev, stream)
    assert len(a) == 21 and a[7:17] == (4, 128, 128, 3, 96, 96, 128, 1, 0, 10) and a[18] == 5
    assert hist["loss"] == [2.0, 2.0] and hist["accuracy"] == [0.25, 0.25] and len(hist["images_per_sec"]) == 2
