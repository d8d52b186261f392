"""On-disk dataset formats behind TFDSDataset / MultiTFDSDataset: IDX,
CIFAR binary batches and .npz, TFDS split strings.  Tiny files are written into tmp_path."""
import gzip
import os
import struct

import numpy as np
import pytest

from zookeeper_b200.core import Field, component, configure
from zookeeper_b200.tf import MultiTFDSDataset, TFDSDataset
from zookeeper_b200.tf import formats


def _write_idx(path, arr, gz=False):
    code = {np.dtype(np.uint8): 0x08, np.dtype(">i4"): 0x0C}[arr.dtype]
    blob = bytes([0, 0, code, arr.ndim]) + struct.pack(">" + "I" * arr.ndim, *arr.shape) + arr.tobytes()
    with (gzip.open(path + ".gz", "wb") if gz else open(path, "wb")) as f:
        f.write(blob)


def _mnist_dir(root, n_train=50, n_test=20, gz=False):
    rng = np.random.default_rng(0)
    d = os.path.join(root, "mnist")
    os.makedirs(d)
    data = {}
    for split, prefix, n in (("train", "train", n_train), ("test", "t10k", n_test)):
        x = rng.integers(0, 256, (n, 28, 28), dtype=np.uint8)
        y = rng.integers(0, 10, (n,), dtype=np.uint8)
        y[0] = 9
        _write_idx(os.path.join(d, f"{prefix}-images-idx3-ubyte"), x, gz)
        _write_idx(os.path.join(d, f"{prefix}-labels-idx1-ubyte"), y, gz)
        data[split] = (x, y)
    return data


def test_idx_reader_roundtrip_and_errors(tmp_path):
    a = np.arange(2 * 3 * 4, dtype=np.uint8).reshape(2, 3, 4)
    _write_idx(str(tmp_path / "a.idx"), a)
    assert np.array_equal(formats.read_idx(str(tmp_path / "a.idx")), a)
    b = np.arange(-3, 3, dtype=">i4")
    _write_idx(str(tmp_path / "b.idx"), b, gz=True)
    assert np.array_equal(formats.read_idx(str(tmp_path / "b.idx.gz")), b)
    (tmp_path / "bad").write_bytes(b"\x01\x02\x03\x04rest")
    with pytest.raises(ValueError, match="not an IDX file"):
        formats.read_idx(str(tmp_path / "bad"))
    blob = (tmp_path / "a.idx").read_bytes()
    (tmp_path / "short").write_bytes(blob[:-5])
    with pytest.raises(ValueError, match="truncated"):
        formats.read_idx(str(tmp_path / "short"))


@pytest.mark.parametrize("gz", [False, True])
def test_tfds_dataset_reads_mnist_idx_with_split_strings(tmp_path, gz):
    data = _mnist_dir(str(tmp_path), gz=gz)

    @component
    class Mnist(TFDSDataset):
        name = Field("mnist:3.0.1")
        train_split = Field("train[:80%]")
        validation_split = Field("train[80%:]")
        test_split = Field("test")

    d = Mnist()
    configure(d, {"data_dir": str(tmp_path)})
    plan, n = d.train()
    assert n == 40 and plan.images.shape == (40, 28, 28, 1) and plan.images.dtype == np.uint8
    assert np.array_equal(plan.images[..., 0], data["train"][0][:40])
    assert np.array_equal(plan.labels, data["train"][1][:40].astype(np.int64))
    vplan, vn = d.validation()
    assert vn == 10 and np.array_equal(vplan.images[..., 0], data["train"][0][40:])
    tplan, tn = d.test()
    assert tn == 20 and np.array_equal(tplan.labels, data["test"][1].astype(np.int64))
    assert d.num_classes == 10
    assert d.num_examples("train+test") == 70 and d.num_examples("train[5:15]+test[-5:]") == 15


def test_tfds_dataset_missing_data_warns_and_raises(tmp_path, capsys):
    @component
    class Nope(TFDSDataset):
        name = Field("not_there")
        train_split = Field("train")

    d = Nope()
    configure(d, {"data_dir": str(tmp_path)})
    with pytest.raises(FileNotFoundError, match="not found under"):
        d.train()
    assert "Field 'download' of component Nope is False" in capsys.readouterr().err
    with pytest.raises(ValueError, match="is not configured with a validation split"):
        d.validation()
    with pytest.raises(ValueError, match="Unable to determine the number of classes"):
        d.num_classes


def test_cifar10_binary_batches_and_npz(tmp_path):
    rng = np.random.default_rng(1)
    d = tmp_path / "cifar10"
    d.mkdir()
    recs = {}
    for name, n in [(f"data_batch_{i}.bin", 6) for i in range(1, 6)] + [("test_batch.bin", 4)]:
        lab = rng.integers(0, 10, (n, 1), dtype=np.uint8)
        chw = rng.integers(0, 256, (n, 3, 32, 32), dtype=np.uint8)
        np.concatenate([lab, chw.reshape(n, -1)], axis=1).tofile(str(d / name))
        recs[name] = (lab[:, 0], chw)
    x, y = formats.load_cifar_bin(str(d), "train")
    assert x.shape == (30, 32, 32, 3) and y.shape == (30,)
    assert np.array_equal(x[6:12], recs["data_batch_2.bin"][1].transpose(0, 2, 3, 1))
    assert np.array_equal(y[6:12], recs["data_batch_2.bin"][0])
    with pytest.raises(ValueError, match="splits 'train' and 'test'"):
        formats.load_cifar_bin(str(d), "validation")

    np.savez(str(tmp_path / "other.npz"), x_train=rng.integers(0, 256, (8, 32, 32, 3), dtype=np.uint8),
             y_train=np.arange(8) % 4, x_test=rng.integers(0, 256, (3, 32, 32, 3), dtype=np.uint8), y_test=np.arange(3))

    @component
    class Both(MultiTFDSDataset):
        train_split = Field(lambda: {"cifar10": "train[:50%]", "other": "train"})
        validation_split = Field(lambda: {"cifar10": "test"})

    m = Both()
    configure(m, {"data_dir": str(tmp_path)})
    plan, n = m.train()
    assert n == 23 and plan.images.shape == (23, 32, 32, 3)
    assert np.array_equal(plan.labels[15:], np.arange(8) % 4)
    assert m.validation()[1] == 4
    with pytest.raises(ValueError, match="is not configured with a test split"):
        m.test()
    assert m.num_classes == 10


def test_split_parser():
    assert formats.base_splits("train+validation") == ["train", "validation"]
    assert formats.base_splits("train") == ["train"]
    assert formats.parse_split("train[:80%]+test[10:]") == [("train", None, "80%"), ("test", "10", None)]
    x = np.arange(10)
    assert formats.take_split(x, x, None, "25%")[0].tolist() == [0, 1]          # round(2.5) -> 2 (to even)
    assert formats.take_split(x, x, "-3", None)[0].tolist() == [7, 8, 9]
    assert formats.take_split(x, x, "8", "4")[0].size == 0
    with pytest.raises(ValueError, match="Unrecognised split"):
        formats.parse_split("train[1:2:3]")
    with pytest.raises(ValueError, match="outside"):
        formats.take_split(x, x, "150%", None)
