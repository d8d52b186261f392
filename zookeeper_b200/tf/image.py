"""Symbolic image ops for ``Preprocessing.input``.

In the reference, ``input()`` is traced once by ``tf.data`` and the resulting TF graph runs per
example on the CPU (examples/larq_experiment.py:24-34).  Here ``input()`` runs once on a symbolic
:class:`ImageSpec`; the recorded chain is lowered to the arguments of ONE fused CUDA kernel
(``zkb_preprocess_u8``) that processes a whole batch.  The functions below mirror the names of the
``tf.image`` / ``tf.cast`` calls the reference example makes.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple


@dataclass
class FusedPreprocess:
    """Arguments of the fused kernel that a recorded op chain lowers to."""

    out_hw: Tuple[int, int]
    pad_size: int = 0          # > 0: centre pad to pad_size then random-crop out_hw
    random_crop: bool = False
    random_flip: bool = False
    scale_div: float = 127.5   # out = float(x) / scale_div - bias_sub
    bias_sub: float = 1.0
    binarize: bool = False


class ImageSpec:
    """A uint8 NHWC image batch flowing through symbolic preprocessing ops."""

    def __init__(self, shape: Tuple[int, int, int], ops: Optional[List[tuple]] = None, dtype: str = "uint8"):
        self.shape = tuple(shape)
        self.ops = list(ops or [])
        self.dtype = dtype

    def _then(self, op: tuple, shape=None, dtype=None) -> "ImageSpec":
        return ImageSpec(shape or self.shape, self.ops + [op], dtype or self.dtype)

    # float32(x) / a - b  (examples/larq_experiment.py:34)
    def __truediv__(self, divisor: float) -> "ImageSpec":
        if self.dtype != "float32":
            raise TypeError("cast the image to float32 before dividing")
        return self._then(("div", float(divisor)))

    def __sub__(self, value: float) -> "ImageSpec":
        if self.dtype != "float32":
            raise TypeError("cast the image to float32 before subtracting")
        return self._then(("sub", float(value)))

    def __repr__(self):
        return f"ImageSpec(shape={self.shape}, dtype={self.dtype}, ops={[o[0] for o in self.ops]})"


def resize_with_crop_or_pad(image: ImageSpec, target_height: int, target_width: int) -> ImageSpec:
    """Centre crop / zero-pad to the target size (``tf.image.resize_with_crop_or_pad``)."""
    return image._then(("crop_or_pad", int(target_height), int(target_width)),
                       shape=(int(target_height), int(target_width), image.shape[2]))


def random_crop(image: ImageSpec, size: Sequence[int]) -> ImageSpec:
    """Uniformly random crop (``tf.image.random_crop``)."""
    size = tuple(int(s) for s in size)
    if len(size) != 3 or size[2] != image.shape[2]:
        raise ValueError(f"random_crop size {size} must be (h, w, {image.shape[2]})")
    if size[0] > image.shape[0] or size[1] > image.shape[1]:
        raise ValueError("random_crop size exceeds the image")
    return image._then(("random_crop", size[0], size[1]), shape=size)


def random_flip_left_right(image: ImageSpec) -> ImageSpec:
    """Flip along W with probability 1/2 (``tf.image.random_flip_left_right``)."""
    return image._then(("random_flip",))


def cast(image: ImageSpec, dtype: str = "float32") -> ImageSpec:
    if str(dtype) not in ("float32", "tf.float32", "<dtype: 'float32'>"):
        raise NotImplementedError("only casts to float32 are supported by the fused kernel")
    return image._then(("cast",), dtype="float32")


def sign(image: ImageSpec) -> ImageSpec:
    """Optional sign-binarisation of the normalised image (north_star item 1; off by default)."""
    return image._then(("sign",))


def lower(image: ImageSpec, source_hw: Tuple[int, int]) -> FusedPreprocess:
    """Lower a recorded chain to the fused kernel, or explain why it cannot be."""
    ops = list(image.ops)
    spec = FusedPreprocess(out_hw=image.shape[:2])
    i = 0
    if i < len(ops) and ops[i][0] == "crop_or_pad":
        _, th, tw = ops[i]
        i += 1
        if i < len(ops) and ops[i][0] == "random_crop":
            if th != tw:
                raise NotImplementedError("padding before a random crop must be square")
            spec.pad_size, spec.random_crop = th, True
            i += 1
        elif (th, tw) != tuple(image.shape[:2]):
            raise NotImplementedError("a centre crop/pad must produce the final spatial size")
    elif i < len(ops) and ops[i][0] == "random_crop":
        if source_hw[0] != source_hw[1]:
            raise NotImplementedError("random crop without padding needs square source images")
        spec.pad_size, spec.random_crop = source_hw[0], True
        i += 1
    if i < len(ops) and ops[i][0] == "random_flip":
        spec.random_flip = True
        i += 1
    if i >= len(ops) or ops[i][0] != "cast":
        raise NotImplementedError(f"expected a float32 cast at op {i} of {[o[0] for o in ops]}")
    i += 1
    if i < len(ops) and ops[i][0] == "div":
        spec.scale_div = ops[i][1]
        i += 1
    else:
        spec.scale_div = 1.0
    if i < len(ops) and ops[i][0] == "sub":
        spec.bias_sub = ops[i][1]
        i += 1
    else:
        spec.bias_sub = 0.0
    if i < len(ops) and ops[i][0] == "sign":
        spec.binarize = True
        i += 1
    if i != len(ops):
        raise NotImplementedError(f"op '{ops[i][0]}' at position {i} is not supported by the fused kernel")
    if (spec.scale_div, spec.bias_sub) != (127.5, 1.0):
        raise NotImplementedError("the fused kernel normalises with x / 127.5 - 1.0 only")
    if spec.random_flip and not spec.random_crop:
        raise NotImplementedError("random flip is fused with the random crop")
    return spec
