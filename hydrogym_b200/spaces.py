"""Minimal ``Box`` space (gymnax ``spaces.Box`` / gymnasium ``spaces.Box`` look-alike).

gymnax and gymnasium are not installable in this image, so the env API is duck-typed;
``Box`` keeps the attributes callers read: ``low``, ``high``, ``shape``, ``dtype``.
Reference use: hydrogym/jax/envs/kolmogorov.py:371-384.
"""
from __future__ import annotations

import numpy as np


class Box:
    def __init__(self, low, high, shape, dtype=np.float32):
        self.low, self.high, self.shape, self.dtype = low, high, tuple(shape), dtype

    def sample(self, rng=None):
        rng = rng if rng is not None else np.random.default_rng()
        lo = np.broadcast_to(np.where(np.isfinite(self.low), self.low, -1.0), self.shape)
        hi = np.broadcast_to(np.where(np.isfinite(self.high), self.high, 1.0), self.shape)
        return rng.uniform(lo, hi).astype(self.dtype)

    def contains(self, x):
        x = np.asarray(x)
        return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

    def __repr__(self):
        return f"Box({self.low}, {self.high}, {self.shape}, {np.dtype(self.dtype).name})"
