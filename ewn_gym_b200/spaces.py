"""Observation / action spaces (envs/ewn.py:61-69).  gymnasium is used when importable; otherwise minimal stand-ins
with the attributes SB3 and the reference touch (shape, dtype, nvec, n, start, sample, seed, contains)."""
from __future__ import annotations

import numpy as np

try:  # pragma: no cover - gymnasium is not installed in the build image
    from gymnasium.spaces import Box, Dict, Discrete, MultiDiscrete  # type: ignore
    HAVE_GYMNASIUM = True
except Exception:
    HAVE_GYMNASIUM = False

    class _Space:
        def __init__(self, shape=None, dtype=None):
            self.shape, self.dtype = shape, None if dtype is None else np.dtype(dtype)
            self._rng = np.random.default_rng()

        def seed(self, seed=None):
            self._rng = np.random.default_rng(seed)
            return [seed]

    class MultiDiscrete(_Space):
        def __init__(self, nvec, dtype=np.int64):
            self.nvec = np.asarray(nvec, dtype=dtype)
            super().__init__(self.nvec.shape, dtype)

        def sample(self):
            return (self._rng.random(self.nvec.shape) * self.nvec).astype(self.dtype)

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(((0 <= x) & (x < self.nvec)).all())

    class Discrete(_Space):
        def __init__(self, n, start=0):
            super().__init__((), np.int64)
            self.n, self.start = int(n), int(start)

        def sample(self):
            return int(self.start + self._rng.integers(self.n))

        def contains(self, x):
            return self.start <= int(x) < self.start + self.n

    class Box(_Space):
        def __init__(self, low, high, shape=None, dtype=np.float32):
            super().__init__(tuple(shape), dtype)
            self.low = np.full(self.shape, low, dtype=dtype)
            self.high = np.full(self.shape, high, dtype=dtype)

        def sample(self):
            return self._rng.integers(self.low, self.high + 1).astype(self.dtype)

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(((self.low <= x) & (x <= self.high)).all())

    class Dict(_Space):
        def __init__(self, spaces=None):
            super().__init__()
            self.spaces = dict(spaces or {})

        def __getitem__(self, k):
            return self.spaces[k]

        def keys(self):
            return self.spaces.keys()

        def items(self):
            return self.spaces.items()

        def sample(self):
            return {k: s.sample() for k, s in self.spaces.items()}

        def seed(self, seed=None):
            return [s.seed(seed) for s in self.spaces.values()]

        def contains(self, x):
            return all(k in x and s.contains(x[k]) for k, s in self.spaces.items())
