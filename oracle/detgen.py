"""Deterministic, library-independent test data -- TEST INFRASTRUCTURE ONLY.

A splitmix64 hash of the element index, mapped to float32-representable values, so that the
golden-vector generator (run once in the build container, next to the reference) and the tests
(run anywhere) build bit-identical inputs without storing them.
"""
from __future__ import annotations

import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = x
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def det_uniform(shape, salt: int, lo: float = -1.0, hi: float = 1.0) -> np.ndarray:
    """float64 array of float32-representable values, uniform in [lo, hi)."""
    n = int(np.prod(shape))
    with np.errstate(over="ignore"):
        idx = np.arange(n, dtype=np.uint64) + np.uint64(salt) * np.uint64(0x1000003)
    h = _splitmix64(idx)
    u = (h >> np.uint64(40)).astype(np.float64) / float(1 << 24)      # 24 random bits
    v = (lo + (hi - lo) * u).astype(np.float32)
    return v.astype(np.float64).reshape(shape)


def det_normalish(shape, salt: int) -> np.ndarray:
    """Sum of four uniforms, variance 1 -- a cheap bell-shaped stand-in for randn."""
    s = sum(det_uniform(shape, salt * 4 + k) for k in range(4))
    return (s * np.sqrt(3.0 / 4.0)).astype(np.float32).astype(np.float64)


def det_labels(n: int, classes: int, salt: int) -> np.ndarray:
    with np.errstate(over="ignore"):
        idx = np.arange(n, dtype=np.uint64) + np.uint64(salt) * np.uint64(0x1000003)
    return (_splitmix64(idx) % np.uint64(classes)).astype(np.int64)


def det_params(sizes, has_bias: bool, salt: int, scale: str = "xavier"):
    """Flat [W0,(b0),W1,(b1)...] with W_l ~ U(-1,1)*sqrt(3/in) (unit-variance/in) -- float32-exact."""
    out = []
    for l in range(len(sizes) - 1):
        fan_in, fan_out = sizes[l], sizes[l + 1]
        w = det_uniform((fan_in, fan_out), salt * 64 + 2 * l) * np.sqrt(3.0 / fan_in)
        out.append(w.astype(np.float32).astype(np.float64))
        if has_bias:
            b = det_uniform((fan_out,), salt * 64 + 2 * l + 1) * 0.1
            out.append(b.astype(np.float32).astype(np.float64))
    return out
