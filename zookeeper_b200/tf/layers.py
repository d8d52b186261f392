"""Layer descriptions with the constructor arguments the reference example passes to
``lq.layers`` / ``tf.keras.layers`` (examples/larq_experiment.py:59-103).  They hold no tensors;
:class:`zookeeper_b200.tf.models.Sequential` lowers a list of them to the engine's blocks."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple, Union


def _pair(v) -> Tuple[int, int]:
    return (int(v), int(v)) if isinstance(v, int) else (int(v[0]), int(v[1]))


@dataclass
class QuantConv2D:
    filters: int
    kernel_size: Union[int, Tuple[int, int]]
    padding: str = "valid"
    input_quantizer: Optional[str] = None
    kernel_quantizer: Optional[str] = None
    kernel_constraint: Optional[str] = None
    use_bias: bool = True
    input_shape: Optional[Tuple[int, int, int]] = None
    strides: Union[int, Tuple[int, int]] = 1

    def __post_init__(self):
        self.kernel_size = _pair(self.kernel_size)
        self.strides = _pair(self.strides)


@dataclass
class QuantDense:
    units: int
    input_quantizer: Optional[str] = None
    kernel_quantizer: Optional[str] = None
    kernel_constraint: Optional[str] = None
    use_bias: bool = True


@dataclass
class BatchNormalization:
    scale: bool = True
    center: bool = True
    momentum: float = 0.99
    epsilon: float = 1e-3


@dataclass
class MaxPool2D:
    pool_size: Union[int, Tuple[int, int]] = (2, 2)
    strides: Optional[Union[int, Tuple[int, int]]] = None

    def __post_init__(self):
        self.pool_size = _pair(self.pool_size)
        self.strides = self.pool_size if self.strides is None else _pair(self.strides)


@dataclass
class Flatten:
    pass


@dataclass
class Activation:
    activation: str
