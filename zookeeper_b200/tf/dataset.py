"""``Dataset`` component contracts (reference zookeeper/tf/dataset.py:11-255) with array payloads.

``train()/validation()/test()`` return ``(DataPlan, num_examples)``: the plan plays the role of the
``tf.data.Dataset`` and accepts the same ``cache/shuffle/repeat/map/batch`` chain.  TFDS itself
(download, TFRecord IO) is outside the hot path; ``TFDSDataset`` keeps its fields but reads the
standard ``<name>.npz`` array file instead (f.3 of SURVEY.md section 8).
"""

import abc
import os
from typing import Dict, Optional, Tuple

import numpy as np

from ..core import utils
from ..core.field import Field
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


class TFDSDataset(Dataset):
    """Same fields as the reference's TFDS wrapper (zookeeper/tf/dataset.py:69-166).  TFDS is not
    part of this build: the split is read from ``<data_dir>/<name>.npz`` (keys ``x_<split>``,
    ``y_<split>``).  ``download`` cannot be honoured (no network) and only warns."""

    name: str = Field()
    data_dir: Optional[str] = Field(None)
    download: bool = Field(False)
    train_split: str = Field()
    validation_split: Optional[str] = Field(None)
    test_split: Optional[str] = Field(None)

    def _file(self) -> str:
        base = self.data_dir or os.path.join(os.path.expanduser("~"), "tensorflow_datasets")
        return os.path.join(base, self.name.split(":")[0] + ".npz")

    def _arrays(self) -> Dict[str, np.ndarray]:
        if not hasattr(self, "_cache"):
            path = self._file()
            if not os.path.exists(path):
                if self.download:
                    utils.warn("Field 'download' is True, but this build has no TFDS / network access.")
                raise FileNotFoundError(
                    f"Dataset '{self.name}' not found at '{path}'. TFDS download/IO is out of scope "
                    "for this build; provide an .npz dump or use SyntheticDataset."
                )
            with np.load(path) as z:
                self._cache = {k: z[k] for k in z.files}
        return self._cache

    @property
    def num_classes(self) -> int:
        try:
            arrays = self._arrays()
            labels = np.concatenate([v.reshape(-1) for k, v in arrays.items() if k.startswith("y_")])
            return int(labels.max()) + 1
        except Exception:
            raise ValueError("Unable to determine the number of classes automatically.") from None

    def num_examples(self, split) -> int:
        return sum(int(self._arrays()[f"x_{s}"].shape[0]) for s in split.split("+"))

    def load(self, split, decoders, shuffle) -> DataPlan:
        arrays = self._arrays()
        xs = [arrays[f"x_{s}"] for s in split.split("+")]
        ys = [arrays[f"y_{s}"] for s in split.split("+")]
        x = np.concatenate(xs)
        if x.ndim == 3:
            x = x[..., None]
        y = np.concatenate([v.reshape(-1) for v in ys]).astype(np.int64)
        return DataPlan(np.ascontiguousarray(x, dtype=np.uint8), y, self.num_classes)

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
