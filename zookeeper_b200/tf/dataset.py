"""``Dataset`` component contracts (reference zookeeper/tf/dataset.py:11-255) with array payloads.

``train()/validation()/test()`` return ``(DataPlan, num_examples)``: the plan plays the role of the
``tf.data.Dataset`` and accepts the same ``cache/shuffle/repeat/map/batch`` chain.  TFDS itself
(download, TFRecord IO) is outside the hot path; ``TFDSDataset`` / ``MultiTFDSDataset`` keep their
fields but read the formats the datasets are published in -- IDX, CIFAR binary batches, ``.npz``
(``formats.py``; f.3 of SURVEY.md section 8).
"""

import abc
import os
from typing import Dict, Optional, Tuple

import numpy as np

from ..core import utils
from ..core.field import Field
from . import formats
from .data import DataPlan


class Dataset(abc.ABC):
    """Abstract dataset: concrete sub-classes implement ``train`` and optionally ``validation``
    / ``test`` (reference zookeeper/tf/dataset.py:11-46)."""

    @abc.abstractmethod
    def train(self, decoders=None) -> Tuple[DataPlan, int]:
        """The training data and the number of training examples."""
        raise NotImplementedError

    def validation(self, decoders=None) -> Tuple[DataPlan, int]:
        raise ValueError(
            f"Dataset '{self.__class__.__name__}' is not configured with validation "
            "data."
        )

    def test(self, decoders=None) -> Tuple[DataPlan, int]:
        raise ValueError(f"Dataset '{self.__class__.__name__}' is not configured with test data.")


class SyntheticDataset(Dataset):
    """i.i.d. uniform uint8 images and uniform labels (SURVEY.md section 8d "synthetic inputs")."""

    image_shape: Tuple[int, int, int] = Field((28, 28, 1))
    num_classes: int = Field(10)
    num_train_examples: int = Field(60000)
    num_validation_examples: int = Field(10000)
    seed: int = Field(1234)

    def _make(self, n: int, salt: int) -> DataPlan:
        rng = np.random.default_rng(self.seed + salt)
        images = rng.integers(0, 256, (n,) + tuple(self.image_shape), dtype=np.uint8)
        labels = rng.integers(0, self.num_classes, (n,)).astype(np.int64)
        return DataPlan(images, labels, self.num_classes, seed=self.seed + salt)

    def train(self, decoders=None) -> Tuple[DataPlan, int]:
        return self._make(self.num_train_examples, 0), self.num_train_examples

    def validation(self, decoders=None) -> Tuple[DataPlan, int]:
        if self.num_validation_examples <= 0:
            return super().validation(decoders)
        return self._make(self.num_validation_examples, 1), self.num_validation_examples


class ArrayDataset(Dataset):
    """Images/labels from an ``.npz`` file with keys ``x_train, y_train, x_test, y_test`` (the
    layout of keras' ``mnist.npz`` / ``cifar10`` dumps); images uint8 NHW or NHWC."""

    path: str = Field()
    num_classes: int = Field(10)

    def _load(self, split: str) -> Tuple[DataPlan, int]:
        if not os.path.exists(self.path):
            raise FileNotFoundError(f"Dataset file '{self.path}' does not exist.")
        with np.load(self.path) as z:
            x, y = z[f"x_{split}"], z[f"y_{split}"]
        if x.ndim == 3:
            x = x[..., None]
        x = np.ascontiguousarray(x, dtype=np.uint8)
        y = np.ascontiguousarray(y.reshape(-1), dtype=np.int64)
        return DataPlan(x, y, self.num_classes), int(x.shape[0])

    def train(self, decoders=None) -> Tuple[DataPlan, int]:
        return self._load("train")

    def validation(self, decoders=None) -> Tuple[DataPlan, int]:
        return self._load("test")


class _Store:
    """Resolves a TFDS-style dataset name to arrays on disk.  Looked up under ``data_dir`` (default
    ``~/tensorflow_datasets``), in this order: ``<name>.npz`` (keys ``x_<split>`` / ``y_<split>``),
    a directory ``<name>/`` holding IDX files (MNIST family), CIFAR-10 binary batches or the
    CIFAR-100 ``train.bin`` / ``test.bin`` pair.  A ``:version`` suffix of the name is ignored."""

    def __init__(self, data_dir: Optional[str]):
        self.base = data_dir or os.path.join(os.path.expanduser("~"), "tensorflow_datasets")
        self._cache: Dict[Tuple[str, str], Tuple[np.ndarray, np.ndarray]] = {}

    def _where(self, name: str) -> Tuple[str, str]:
        stem = name.split(":")[0]
        npz = os.path.join(self.base, stem + ".npz")
        if os.path.exists(npz):
            return "npz", npz
        d = os.path.join(self.base, stem)
        if os.path.isdir(d):
            files = set(os.listdir(d))
            if any("idx3-ubyte" in f for f in files):
                return "idx", d
            if "data_batch_1.bin" in files or "test_batch.bin" in files:
                return "cifar10", d
            if "train.bin" in files or "test.bin" in files:
                return "cifar100", d
        raise FileNotFoundError(
            f"Dataset '{name}' not found under '{self.base}' (looked for '{stem}.npz' and a directory "
            f"'{stem}/' with IDX or CIFAR binary files). TFDS download/TFRecord IO is not part of this "
            "build; provide one of these formats or use SyntheticDataset."
        )

    def base_split(self, name: str, split: str) -> Tuple[np.ndarray, np.ndarray]:
        key = (name, split)
        if key not in self._cache:
            kind, path = self._where(name)
            if kind == "npz":
                arrays = formats.load_npz(path)
                if f"x_{split}" not in arrays:
                    have = sorted(k[2:] for k in arrays if k.startswith("x_"))
                    raise ValueError(f"Dataset '{name}' has no split '{split}' (available: {have}).")
                x, y = arrays[f"x_{split}"], arrays[f"y_{split}"]
            elif kind == "idx":
                x, y = formats.load_idx_pair(path, split)
            elif kind == "cifar10":
                x, y = formats.load_cifar_bin(path, split)
            else:
                x, y = formats.load_cifar_bin(path, split, label_bytes=2)
            if x.ndim == 3:
                x = x[..., None]
            self._cache[key] = (np.ascontiguousarray(x, dtype=np.uint8),
                                np.ascontiguousarray(np.asarray(y).reshape(-1), dtype=np.int64))
        return self._cache[key]

    def split(self, name: str, split: str) -> Tuple[np.ndarray, np.ndarray]:
        parts = [formats.take_split(*self.base_split(name, n), lo, hi) for n, lo, hi in formats.parse_split(split)]
        if len(parts) == 1:
            return parts[0]
        return np.concatenate([p[0] for p in parts]), np.concatenate([p[1] for p in parts])

    def num_classes(self, name: str) -> int:
        best = 0
        for split in ("train", "test", "validation"):
            try:
                best = max(best, int(self.base_split(name, split)[1].max()) + 1)
            except (ValueError, FileNotFoundError):
                continue
        if best == 0:
            raise ValueError("Unable to determine the number of classes automatically.")
        return best


def _warn_download(component, download: bool):
    if not download:
        utils.warn(
            f"Field 'download' of component {component.__class__.__name__} is "
            "False. If the TFDS dataset is not downloaded, set 'download' "
            "to True to call `download_and_prepare()` automatically."
        )
    else:
        utils.warn("Field 'download' is True, but this build has no TFDS / network access.")


class TFDSDataset(Dataset):
    """Same fields as the reference's TFDS wrapper (zookeeper/tf/dataset.py:69-166).  TFDS is not
    part of this build: the named dataset is read from the formats :class:`_Store` lists, with
    TFDS split strings (``"train"``, ``"train[:90%]"``, ``"train+test"``).  ``download`` cannot be
    honoured (no network) and only warns."""

    name: str = Field()
    data_dir: Optional[str] = Field(None)
    download: bool = Field(False)
    train_split: str = Field()
    validation_split: Optional[str] = Field(None)
    test_split: Optional[str] = Field(None)

    def _store(self) -> _Store:
        if not hasattr(self, "_store_obj"):
            object.__setattr__(self, "_store_obj", _Store(self.data_dir))
        return self._store_obj

    @property
    def num_classes(self) -> int:
        try:
            return self._store().num_classes(self.name)
        except Exception:
            raise ValueError("Unable to determine the number of classes automatically.") from None

    def num_examples(self, split) -> int:
        return int(self._store().split(self.name, split)[0].shape[0])

    def load(self, split, decoders, shuffle) -> DataPlan:
        try:
            x, y = self._store().split(self.name, split)
        except FileNotFoundError:
            _warn_download(self, self.download)
            raise
        return DataPlan(x, y, self.num_classes)

    def train(self, decoders=None) -> Tuple[DataPlan, int]:
        return self.load(self.train_split, decoders, True), self.num_examples(self.train_split)

    def validation(self, decoders=None) -> Tuple[DataPlan, int]:
        if self.validation_split is None:
            raise ValueError(
                f"Dataset {self.__class__.__name__} is not configured with a "
                "validation split."
            )
        return self.load(self.validation_split, decoders, False), self.num_examples(self.validation_split)

    def test(self, decoders=None) -> Tuple[DataPlan, int]:
        if self.test_split is None:
            raise ValueError(
                f"Dataset {self.__class__.__name__} is not configured with a "
                "test split."
            )
        return self.load(self.test_split, decoders, False), self.num_examples(self.test_split)


class MultiTFDSDataset(Dataset):
    """Several named datasets concatenated (zookeeper/tf/dataset.py:169-255): every ``*_split`` is
    a mapping from dataset name to split string; an empty mapping means the split does not exist."""

    data_dir: Optional[str] = Field(None)
    download: bool = Field(False)
    train_split: Dict[str, str] = Field()
    validation_split: Dict[str, str] = Field(lambda: {})
    test_split: Dict[str, str] = Field(lambda: {})

    def _store(self) -> _Store:
        if not hasattr(self, "_store_obj"):
            object.__setattr__(self, "_store_obj", _Store(self.data_dir))
        return self._store_obj

    @property
    def num_classes(self) -> int:
        return max(self._store().num_classes(name) for name in self.train_split)

    def num_examples(self, splits) -> int:
        return sum(int(self._store().split(name, split)[0].shape[0]) for name, split in splits.items())

    def load(self, splits, decoders, shuffle) -> DataPlan:
        xs, ys = [], []
        for name, split in splits.items():
            try:
                x, y = self._store().split(name, split)
            except FileNotFoundError:
                _warn_download(self, self.download)
                raise
            xs.append(x)
            ys.append(y)
        shapes = {x.shape[1:] for x in xs}
        if len(shapes) != 1:
            raise ValueError(f"Datasets {list(splits)} have different image shapes {sorted(shapes)}.")
        return DataPlan(np.concatenate(xs), np.concatenate(ys), self.num_classes)

    def train(self, decoders=None) -> Tuple[DataPlan, int]:
        return self.load(self.train_split, decoders, True), self.num_examples(self.train_split)

    def validation(self, decoders=None) -> Tuple[DataPlan, int]:
        if not self.validation_split:
            raise ValueError(
                f"Dataset {self.__class__.__name__} is not configured with a "
                "validation split."
            )
        return self.load(self.validation_split, decoders, False), self.num_examples(self.validation_split)

    def test(self, decoders=None) -> Tuple[DataPlan, int]:
        if not self.test_split:
            raise ValueError(
                f"Dataset {self.__class__.__name__} is not configured with a "
                "test split."
            )
        return self.load(self.test_split, decoders, False), self.num_examples(self.test_split)
