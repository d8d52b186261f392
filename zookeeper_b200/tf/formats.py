"""On-disk image dataset formats behind the ``Dataset`` components.

The reference reads everything through TensorFlow Datasets (zookeeper/tf/dataset.py:93, 123-130,
201, 210-217), i.e. TFRecord shards prepared by ``download_and_prepare``.  TFDS is not part of this
build; the loaders below read the formats the same datasets are published in and return plain
uint8 NHWC / int64 arrays, which is what the device-resident pool of the engine is filled from.

* IDX (``train-images-idx3-ubyte[.gz]`` ...): MNIST, Fashion-MNIST, EMNIST, KMNIST.
* CIFAR binary batches (``data_batch_1.bin`` ...): one label byte (two for CIFAR-100) followed by
  3072 bytes in channel-major order per record.
* ``.npz`` archives with ``x_<split>`` / ``y_<split>`` arrays (the Keras dumps).
"""

import gzip
import os
import re
import struct
from typing import Dict, List, Optional, Tuple

import numpy as np

_IDX_DTYPES = {0x08: np.uint8, 0x09: np.int8, 0x0B: ">i2", 0x0C: ">i4", 0x0D: ">f4", 0x0E: ">f8"}


def _open(path: str):
    return gzip.open(path, "rb") if path.endswith(".gz") else open(path, "rb")


def _first_existing(directory: str, names: List[str]) -> str:
    for n in names:
        for cand in (n, n + ".gz"):
            p = os.path.join(directory, cand)
            if os.path.exists(p):
                return p
    raise FileNotFoundError(f"None of {names} (optionally .gz) found in '{directory}'.")


def read_idx(path: str) -> np.ndarray:
    """One IDX file: magic ``00 00 <dtype> <ndim>``, ``ndim`` big-endian uint32 sizes, then the data."""
    with _open(path) as f:
        head = f.read(4)
        if len(head) != 4 or head[0] != 0 or head[1] != 0 or head[2] not in _IDX_DTYPES:
            raise ValueError(f"'{path}' is not an IDX file (bad magic {head!r}).")
        ndim = head[3]
        dims = struct.unpack(">" + "I" * ndim, f.read(4 * ndim))
        dtype = np.dtype(_IDX_DTYPES[head[2]])
        want = int(np.prod(dims, dtype=np.int64)) * dtype.itemsize
        raw = f.read(want)
        if len(raw) != want:
            raise ValueError(f"'{path}' is truncated: expected {want} data bytes, found {len(raw)}.")
    return np.frombuffer(raw, dtype=dtype).reshape(dims)


def load_idx_pair(directory: str, split: str) -> Tuple[np.ndarray, np.ndarray]:
    """(images uint8 [N,H,W,1], labels int64 [N]) of an MNIST-style directory; split is ``train`` or
    ``test`` (file prefix ``t10k``)."""
    prefix = {"train": "train", "test": "t10k"}.get(split)
    if prefix is None:
        raise ValueError(f"IDX datasets have the splits 'train' and 'test', not '{split}'.")
    x = read_idx(_first_existing(directory, [f"{prefix}-images-idx3-ubyte", f"{prefix}-images.idx3-ubyte"]))
    y = read_idx(_first_existing(directory, [f"{prefix}-labels-idx1-ubyte", f"{prefix}-labels.idx1-ubyte"]))
    if x.ndim != 3 or y.ndim != 1 or x.shape[0] != y.shape[0]:
        raise ValueError(f"IDX image/label files in '{directory}' do not match: {x.shape} vs {y.shape}.")
    return np.ascontiguousarray(x[..., None], dtype=np.uint8), y.astype(np.int64)


def load_cifar_bin(directory: str, split: str, label_bytes: int = 1, fine: bool = True) -> Tuple[np.ndarray, np.ndarray]:
    """(images uint8 [N,32,32,3], labels int64 [N]) from the CIFAR binary version.  CIFAR-10:
    ``data_batch_{1..5}.bin`` / ``test_batch.bin`` with one label byte; CIFAR-100: ``train.bin`` /
    ``test.bin`` with a coarse and a fine label byte (``label_bytes=2``)."""
    if split not in ("train", "test"):
        raise ValueError(f"CIFAR has the splits 'train' and 'test', not '{split}'.")
    if label_bytes == 1:
        names = [f"data_batch_{i}.bin" for i in range(1, 6)] if split == "train" else ["test_batch.bin"]
    else:
        names = [f"{split}.bin"]
    rec = label_bytes + 3072
    xs, ys = [], []
    for n in names:
        p = os.path.join(directory, n)
        if not os.path.exists(p):
            raise FileNotFoundError(f"CIFAR batch '{p}' does not exist.")
        raw = np.fromfile(p, dtype=np.uint8)
        if raw.size % rec:
            raise ValueError(f"'{p}': size {raw.size} is not a multiple of the {rec}-byte record.")
        raw = raw.reshape(-1, rec)
        ys.append(raw[:, label_bytes - 1 if fine else 0].astype(np.int64))
        xs.append(raw[:, label_bytes:].reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1))
    return np.ascontiguousarray(np.concatenate(xs)), np.concatenate(ys)


def load_npz(path: str) -> Dict[str, np.ndarray]:
    if not os.path.exists(path):
        raise FileNotFoundError(f"Dataset file '{path}' does not exist.")
    with np.load(path) as z:
        return {k: z[k] for k in z.files}


# ---------------------------------------------------------------------------------- split strings
_SLICE = re.compile(r"^\s*([A-Za-z_][A-Za-z0-9_]*)\s*(?:\[\s*(-?\d+%?)?\s*:\s*(-?\d+%?)?\s*\])?\s*$")


def base_splits(split: str) -> List[str]:
    """``"train+validation"`` -> ``["train", "validation"]`` (zookeeper/tf/dataset.py:49-66)."""
    return [s.strip() for s in split.split("+")] if "+" in split else [split]


def _bound(text: Optional[str], n: int, default: int) -> int:
    if text is None:
        return default
    if text.endswith("%"):
        pct = int(text[:-1])
        if not -100 <= pct <= 100:
            raise ValueError(f"Percent slice boundary '{text}' is outside [-100%, 100%].")
        v = int(round(n * pct / 100.0))      # TFDS rounds percent boundaries to the closest example
    else:
        v = int(text)
    if v < 0:
        v += n
    return min(max(v, 0), n)


def parse_split(split: str) -> List[Tuple[str, Optional[str], Optional[str]]]:
    """TFDS split syntax: ``"train"``, ``"train[:80%]"``, ``"train[100:200]"``, joined with ``+``.
    Returns ``[(name, lo_text, hi_text), ...]``."""
    out = []
    for part in base_splits(split):
        m = _SLICE.match(part)
        if not m:
            raise ValueError(f"Unrecognised split specification '{part}'.")
        out.append((m.group(1), m.group(2), m.group(3)))
    return out


def take_split(x: np.ndarray, y: np.ndarray, lo: Optional[str], hi: Optional[str]) -> Tuple[np.ndarray, np.ndarray]:
    n = int(x.shape[0])
    a, b = _bound(lo, n, 0), _bound(hi, n, n)
    if b < a:
        b = a
    return x[a:b], y[a:b]
