"""Deterministic synthetic inputs shared by the golden generator and the tests.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Integer-hash based, so the same arrays come out on any machine and any numpy /
torch version (torch's CPU generator is not guaranteed stable across builds).
"""
from __future__ import annotations

import numpy as np

_M64 = (1 << 64) - 1


def _mix(x: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser on a uint64 array."""
    x = x.copy()
    x ^= x >> np.uint64(30)
    x *= np.uint64(0xBF58476D1CE4E5B9)
    x ^= x >> np.uint64(27)
    x *= np.uint64(0x94D049BB133111EB)
    x ^= x >> np.uint64(31)
    return x


def _stream(numel: int, *key: int) -> np.ndarray:
    k = 0x9E3779B97F4A7C15
    for v in key:
        k = (k * 0x100000001B3 + (int(v) + 1) * 0xD6E8FEB86659FD93) & _M64
    with np.errstate(over="ignore"):
        return _mix(np.arange(numel, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15) + np.uint64(k))


def grid_grad(numel: int, rank: int, step: int, tensor: int) -> np.ndarray:
    """Gradient on a coarse binary grid: an integer in [-4096, 4096) times 2**-20.
    Any sum of up to 2**10 of them, in any order, and its division by a power
    of two, is exact in fp32 - so the averaged gradient is bit-identical for
    every world size and every reduction order."""
    v = (_stream(numel, 1, rank, step, tensor) & np.uint64(0x1FFF)).astype(np.int64) - 4096
    return (v.astype(np.float32) * np.float32(2.0 ** -20)).astype(np.float32)


def dense_grad(numel: int, rank: int, step: int, tensor: int) -> np.ndarray:
    """Full-mantissa gradient, roughly uniform in (-1e-2, 1e-2)."""
    v = (_stream(numel, 2, rank, step, tensor) >> np.uint64(40)).astype(np.float64)  # 24 bits
    return ((v / float(1 << 23) - 1.0) * 1e-2).astype(np.float32)


def init_param(numel: int, tensor: int) -> np.ndarray:
    v = (_stream(numel, 3, tensor) >> np.uint64(40)).astype(np.float64)
    return ((v / float(1 << 23) - 1.0) * 0.5).astype(np.float32)


# ragged tensor set: empty, scalars, odd tails, just under/over vector and
# chunk boundaries, one larger tensor.
RAGGED_NUMELS = (1, 2, 3, 4, 5, 7, 8, 31, 32, 33, 127, 1000, 1023, 1025, 2048, 4099)

SGD_VARIANTS = {
    "A": dict(lr=0.01, momentum=0.9),                                   # codes/task2/model.py:131
    "B": dict(lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=True),
    "C": dict(lr=0.05, momentum=0.8, dampening=0.1, weight_decay=5e-4),
    "D": dict(lr=0.001),                                                # plain GD, momentum 0
}
