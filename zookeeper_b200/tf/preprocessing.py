"""``Preprocessing`` component contract (reference zookeeper/tf/preprocessing.py:11-75)."""

import functools
import inspect
from typing import Any, Dict, Optional, Tuple

from ..core.field import Field


def pass_training_kwarg(function, training=False):
    """Forward ``training`` only to methods that declare it (reference :11-14)."""
    if "training" in inspect.signature(function).parameters:
        return functools.partial(function, training=training)
    return function


class Preprocessing:
    """Wraps per-example preprocessing.  ``input``/``output`` receive a dict of symbolic features
    (``{"image": ImageSpec, "label": "label"}``) and return what the model consumes / is trained
    against; the recorded image ops are lowered to one fused CUDA kernel per batch."""

    decoders: Optional[Dict[str, Any]] = Field(None)

    # The shape of the processed input. Must match the output of `input()`.
    input_shape: Tuple[int, int, int] = Field()

    def input(self, data, training):
        raise NotImplementedError("Must be implemented in subclasses.")

    def output(self, data, training):
        raise NotImplementedError("Must be implemented in subclasses.")

    def __call__(self, data, training=False):
        input_fn = pass_training_kwarg(self.input, training=training)
        output_fn = pass_training_kwarg(self.output, training=training)
        return input_fn(data), output_fn(data)
