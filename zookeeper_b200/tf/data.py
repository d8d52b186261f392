"""Lazy dataset plans: the ``tf.data.Dataset`` call chain of the reference's ``run()``
(``.cache().shuffle(n).repeat().map(fn).batch(bs)``, examples/larq_experiment.py:126-139)
recorded instead of executed, so that the train loop can realise it on the GPU: the cached dataset
is a resident uint8 pool, shuffle becomes an index stream, map+batch become one fused kernel."""
from __future__ import annotations

import functools
import inspect
from dataclasses import dataclass, replace
from typing import Any, Callable, Iterator, Optional

import numpy as np


@dataclass(frozen=True)
class DataPlan:
    images: Any                       # uint8 [n, H, W, C] (numpy or torch)
    labels: Any                       # int64 [n]
    num_classes: int
    cached: bool = False
    shuffle_buffer: int = 0
    repeated: bool = False
    map_fn: Optional[Callable] = None
    batch_size: int = 0
    seed: int = 0

    def __len__(self) -> int:
        return int(self.images.shape[0])

    def cache(self) -> "DataPlan":
        return replace(self, cached=True)

    def shuffle(self, buffer_size: int, seed: Optional[int] = None) -> "DataPlan":
        return replace(self, shuffle_buffer=int(buffer_size), seed=self.seed if seed is None else int(seed))

    def repeat(self) -> "DataPlan":
        return replace(self, repeated=True)

    def map(self, fn: Callable) -> "DataPlan":
        if self.batch_size:
            raise NotImplementedError("map after batch is not supported; map per example, then batch")
        return replace(self, map_fn=fn)

    def batch(self, batch_size: int) -> "DataPlan":
        return replace(self, batch_size=int(batch_size))

    def prefetch(self, *_args, **_kwargs) -> "DataPlan":
        return self  # everything is device resident; nothing to prefetch

    # ------------------------------------------------------------------ realisation helpers
    @property
    def training(self) -> bool:
        """Was the mapped preprocessing bound with ``training=True``?"""
        fn = self.map_fn
        while isinstance(fn, functools.partial):
            if "training" in fn.keywords:
                return bool(fn.keywords["training"])
            fn = fn.func
        return False

    @property
    def preprocessing(self):
        fn = self.map_fn
        while isinstance(fn, functools.partial):
            fn = fn.func
        return fn

    def index_stream(self, epoch: int = 0) -> Iterator[np.ndarray]:
        """Batches of sample indices.  ``shuffle(buffer)`` follows tf.data's streaming shuffle: a
        buffer of ``buffer`` elements is kept full from the sequential stream and each output is
        a uniformly random element of the buffer.  ``.repeat()`` before ``.batch()`` (the order of
        the reference's chain) makes one endless sample stream that is cut into batches, so the
        ``n % batch`` tail of a pass fills the start of the next pass's first batch; without
        ``repeat`` the incomplete last batch is dropped (the engine's batch size is fixed; the
        reference never consumes it either: its step counts are ``n // batch_size``,
        examples/larq_experiment.py:150-152)."""
        n, bs = len(self), self.batch_size
        if bs <= 0:
            raise ValueError("the plan has no batch size; call .batch(n)")
        rng = np.random.default_rng(self.seed + 7919 * epoch)
        carry = np.empty(0, dtype=np.int64)
        while True:
            order = np.arange(n, dtype=np.int64)
            if self.shuffle_buffer > 1:
                order = _streaming_shuffle(n, min(self.shuffle_buffer, n), rng)
            if carry.size:
                order = np.concatenate([carry, order])
            full = (order.size // bs) * bs
            for s in range(0, full, bs):
                yield order[s:s + bs]
            if not self.repeated:
                return
            carry = order[full:]


def _streaming_shuffle(n: int, buffer: int, rng: np.random.Generator) -> np.ndarray:
    buf = np.arange(buffer, dtype=np.int64)
    out = np.empty(n, dtype=np.int64)
    picks = rng.integers(0, buffer, size=n)
    nxt = buffer
    for i in range(n):
        j = picks[i] % max(1, min(buffer, n - i))
        out[i] = buf[j]
        if nxt < n:
            buf[j] = nxt
            nxt += 1
        else:
            # stream exhausted: shrink the buffer by moving its last live element into the hole
            last = min(buffer, n - i) - 1
            buf[j] = buf[last]
    return out
