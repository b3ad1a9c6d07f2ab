"""CPU restatement of DistributedSampler's index shard (integer, bit-exact).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Follows the object the reference constructs at codes/task2/model.py:124
(`DistributedSampler(dataset, num_replicas, rank)`), whose arithmetic lives in
third-party torch: torch/utils/data/distributed.py:66-141 (sizes, pad /
truncate, stride) and ATen's randperm_cpu + at::mt19937 (32-bit draws,
forward Fisher-Yates).  The constructor shape mirrored by the reference's own
skeleton is codes/task3/sampler.py:5-25.

Pinned against torch itself in tests/test_oracle_golden.py and against
tests/golden/sampler_*.npz (written by oracle/make_golden.py).
"""
from __future__ import annotations

import ctypes
import math
import os
import subprocess

import numpy as np

_N, _M = 624, 397
_UPPER, _LOWER, _MATRIX_A = 0x80000000, 0x7FFFFFFF, 0x9908B0DF


def mt19937_init(seed: int) -> np.ndarray:
    """init_genrand: only the low 32 bits of the seed are used (at::mt19937)."""
    st = np.empty(_N, np.uint64)
    st[0] = seed & 0xFFFFFFFF
    for i in range(1, _N):
        prev = int(st[i - 1])
        st[i] = (1812433253 * (prev ^ (prev >> 30)) + i) & 0xFFFFFFFF
    return st.astype(np.uint32)


def _twist_block(st: np.ndarray) -> np.ndarray:
    """One regeneration of the 624-word state, vectorised in the three
    segments inside which no word depends on another of the same segment."""
    st = st.copy()

    def seg(lo, hi, src_lo):
        y = (st[lo:hi] & np.uint32(_UPPER)) | (st[lo + 1:hi + 1] & np.uint32(_LOWER))
        mag = np.where(y & np.uint32(1), np.uint32(_MATRIX_A), np.uint32(0))
        st[lo:hi] = st[src_lo:src_lo + (hi - lo)] ^ (y >> np.uint32(1)) ^ mag

    seg(0, _N - _M, _M)                 # 0..226   reads old 397..623
    seg(_N - _M, 2 * (_N - _M), 0)      # 227..453 reads new 0..226
    seg(2 * (_N - _M), _N - 1, _N - _M)  # 454..622 reads new 227..395
    y = (st[_N - 1] & np.uint32(_UPPER)) | (st[0] & np.uint32(_LOWER))
    st[_N - 1] = st[_M - 1] ^ (y >> np.uint32(1)) ^ (np.uint32(_MATRIX_A) if y & np.uint32(1) else np.uint32(0))
    return st


def _temper(y: np.ndarray) -> np.ndarray:
    y = y ^ (y >> np.uint32(11))
    y = y ^ ((y << np.uint32(7)) & np.uint32(0x9D2C5680))
    y = y ^ ((y << np.uint32(15)) & np.uint32(0xEFC60000))
    return y ^ (y >> np.uint32(18))


def mt19937_draws(seed: int, count: int) -> np.ndarray:
    """First `count` 32-bit outputs of mt19937(seed)."""
    out = np.empty(max(count, 0), np.uint32)
    st = mt19937_init(seed)
    done = 0
    while done < count:
        st = _twist_block(st)
        take = min(_N, count - done)
        out[done:done + take] = _temper(st[:take])
        done += take
    return out


def randperm(n: int, seed: int) -> np.ndarray:
    """torch.randperm(n, generator=manual_seed(seed)) for n < 2**32 / 20:
    for i in 0..n-2: z = draw % (n - i); swap(r[i], r[i + z])."""
    if n >= (1 << 32) // 20:
        raise ValueError("64-bit randperm path is out of scope")
    r = list(range(n))
    draws = mt19937_draws(seed, max(n - 1, 0)).tolist()
    for i in range(n - 1):
        j = i + draws[i] % (n - i)
        r[i], r[j] = r[j], r[i]
    return np.asarray(r, np.int64)


def shard_sizes(n: int, world: int, drop_last: bool) -> tuple[int, int]:
    """(num_samples, total_size): torch/utils/data/distributed.py:93-105."""
    if drop_last and n % world != 0:
        num = math.ceil((n - world) / world)
    else:
        num = math.ceil(n / world)
    return num, num * world


def distributed_sampler_indices(n, world, rank, seed=0, epoch=0, shuffle=True,
                                drop_last=False, perm=None) -> np.ndarray:
    """The list DistributedSampler.__iter__ walks, as int64."""
    if not 0 <= rank < world:
        raise ValueError(f"Invalid rank {rank}, rank should be in the interval [0, {world - 1}]")
    num, total = shard_sizes(n, world, drop_last)
    if perm is None:
        perm = randperm(n, seed + epoch) if shuffle else np.arange(n, dtype=np.int64)
    t = rank + world * np.arange(num, dtype=np.int64)          # positions in the padded list
    if n == 0:
        return np.empty(0, np.int64)
    src = np.where(t < n, t, (t - n) % n)                      # pad = repeat from the head
    return perm[src]


# ---- C twin (oracle/sampler_oracle.c), used for big n and as CPU baseline ----
_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libsampler_oracle.so")


def build_c():
    os.makedirs(os.path.dirname(_SO), exist_ok=True)
    src = os.path.join(_HERE, "sampler_oracle.c")
    if not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-o", _SO, src])
    return _SO


def load_c():
    lib = ctypes.CDLL(build_c())
    lib.oracle_sampler_shard.restype = ctypes.c_int64
    lib.oracle_sampler_shard.argtypes = [ctypes.c_int64, ctypes.c_int, ctypes.c_int,
                                         ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int,
                                         ctypes.c_int, ctypes.c_void_p]
    lib.oracle_randperm.restype = ctypes.c_int
    lib.oracle_randperm.argtypes = [ctypes.c_int64, ctypes.c_uint64, ctypes.c_void_p]
    return lib


def c_distributed_sampler_indices(n, world, rank, seed=0, epoch=0, shuffle=True,
                                  drop_last=False) -> np.ndarray:
    lib = load_c()
    num, _ = shard_sizes(n, world, drop_last)
    out = np.empty(num, np.int64)
    got = lib.oracle_sampler_shard(n, world, rank, seed, epoch, int(shuffle),
                                   int(drop_last), out.ctypes.data)
    if got != num:
        raise RuntimeError(f"oracle_sampler_shard returned {got}, expected {num}")
    return out


def c_randperm(n, seed) -> np.ndarray:
    lib = load_c()
    out = np.empty(n, np.int64)
    if lib.oracle_randperm(n, seed, out.ctypes.data) != 0:
        raise RuntimeError("oracle_randperm failed")
    return out
