#!/usr/bin/env python
"""MT19937 jump-ahead, as groundwork for splitting K4's serial recurrence over several CTAs.

The sampler's draws are the tempered words x[k+624] of the sequence
    x[0..623] = seeded state,   x[k+624] = x[k+397] ^ twist(x[k], x[k+1]),
which one CTA can only produce 227 at a time (csrc/sampler.cu mt_targets): at
n = 1,281,167 that serial part is ~70 % of the kernel.  The recurrence is linear over GF(2), so
every bit of the word sequence satisfies phi(E) x = 0 for one polynomial phi of degree 19937 in
the shift E: k -> k+1.  With g_J(t) = t^J mod phi(t):

    x[k + J] = XOR over { i : coefficient i of g_J is 1 } of x[k + i]        (k >= 1)

so a CTA that is to start at draw D needs only the first 19937 + 624 words (90 steps of the
227-wide recurrence) and the 2.5 KB polynomial g_D to rebuild its own 624-word window - instead of
D / 227 serial steps.  Polynomials depend on the distance only, not on the seed.

This file computes phi (Berlekamp-Massey on one bit of the sequence), g_J, and checks the identity
against the sequential generator; `--header` writes the polynomials for J = c * stride as a C
array.  Nothing in the library uses it yet (round-2 work); tests/test_mt_jump.py pins the maths.

    python tools/mt_jump.py --check
    python tools/mt_jump.py --header csrc_mt_jump.h --stride 131072 --count 15
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

DEG = 19937
_N, _M = 624, 397


def _seeded_state(seed: int) -> np.ndarray:
    st = np.empty(_N, np.uint64)
    st[0] = seed & 0xFFFFFFFF
    for i in range(1, _N):
        prev = int(st[i - 1])
        st[i] = (1812433253 * (prev ^ (prev >> 30)) + i) & 0xFFFFFFFF
    return st.astype(np.uint32)


def _next_block(st: np.ndarray) -> np.ndarray:
    """The next 624 words, in the three segments inside which no word depends on another."""
    st = st.copy()

    def seg(lo, hi, src):
        y = (st[lo:hi] & np.uint32(0x80000000)) | (st[lo + 1:hi + 1] & np.uint32(0x7FFFFFFF))
        st[lo:hi] = st[src:src + hi - lo] ^ (y >> np.uint32(1)) ^ np.where(y & np.uint32(1), np.uint32(0x9908B0DF), np.uint32(0))

    seg(0, _N - _M, _M)
    seg(_N - _M, 2 * (_N - _M), 0)
    seg(2 * (_N - _M), _N - 1, _N - _M)
    y = (int(st[_N - 1]) & 0x80000000) | (int(st[0]) & 0x7FFFFFFF)
    st[_N - 1] = int(st[_M - 1]) ^ (y >> 1) ^ (0x9908B0DF if y & 1 else 0)
    return st


def word_sequence(seed: int, count: int) -> np.ndarray:
    """x[0..count): the seeded state followed by the untempered recurrence."""
    st = _seeded_state(seed)
    blocks = [st]
    while 624 * len(blocks) < count:
        st = _next_block(st)
        blocks.append(st)
    return np.concatenate(blocks)[:count]


def berlekamp_massey(bits) -> tuple[int, int]:
    """(connection polynomial C as an int with bit i = c_i, its length L) of a GF(2) sequence:
    sum_{i=0..L} c_i s[N-i] = 0 for all N >= L."""
    n = len(bits)
    rev = 0                                   # bit (n-1-k) = s_k, so (rev >> (n-1-N)) has bit i = s[N-i]
    for k, b in enumerate(bits):
        if b:
            rev |= 1 << (n - 1 - k)
    C, B, L, m = 1, 1, 0, 1
    for N in range(n):
        if (C & (rev >> (n - 1 - N))).bit_count() & 1:
            T = C
            C ^= B << m
            if 2 * L <= N:
                L, B, m = N + 1 - L, T, 1
            else:
                m += 1
        else:
            m += 1
    return C, L


def characteristic_polynomial(seed: int = 5489) -> int:
    """phi as an int (bit j = coefficient of t^j): sum_j phi_j x[k+j] = 0 for every k >= 1."""
    x = word_sequence(seed, 2 * DEG + 64)
    bits = ((x[1:] >> np.uint32(7)) & np.uint32(1)).tolist()      # any one bit of the words will do
    C, L = berlekamp_massey(bits)
    assert L == DEG, f"minimal polynomial has degree {L}, expected {DEG}"
    phi = 0
    for i in range(L + 1):                                        # reciprocal: phi_j = c_{L-j}
        if (C >> i) & 1:
            phi |= 1 << (L - i)
    return phi


def _reduce(p: int, phi: int) -> int:
    d = p.bit_length() - 1
    while d >= DEG:
        p ^= phi << (d - DEG)
        d = p.bit_length() - 1
    return p


def _mulmod(a: int, b: int, phi: int) -> int:
    acc = 0
    while b:
        low = b & -b
        acc ^= a << (low.bit_length() - 1)
        b ^= low
    return _reduce(acc, phi)


def jump_polynomial(J: int, phi: int) -> int:
    """t^J mod phi(t)."""
    result, base = 1, 2                       # 1 and t
    while J:
        if J & 1:
            result = _mulmod(result, base, phi)
        J >>= 1
        if J:
            base = _mulmod(base, base, phi)
    return result


def jumped_window(x_prefix: np.ndarray, g: int, length: int = 624, first: int = 1) -> np.ndarray:
    """x[J + first .. J + first + length) from x[0 .. 19937 + first + length)."""
    idx = np.array([i for i in range(DEG) if (g >> i) & 1], dtype=np.int64)
    js = np.arange(first, first + length, dtype=np.int64)
    return np.bitwise_xor.reduce(x_prefix[idx[:, None] + js[None, :]], axis=0)


def check(seed=20231106, J=131072) -> None:
    phi = characteristic_polynomial()
    g = jump_polynomial(J, phi)
    x = word_sequence(seed, J + 2048)
    got = jumped_window(x[: DEG + 1 + 700], g, length=700)
    assert np.array_equal(got, x[J + 1: J + 701]), "jump-ahead identity violated"
    print(f"phi: degree {phi.bit_length() - 1}, {phi.bit_count()} terms; g_{J}: {g.bit_count()} terms; identity holds")


def write_header(path: str, stride: int, count: int) -> None:
    phi = characteristic_polynomial()
    words = (DEG + 31) // 32
    with open(path, "w") as f:
        f.write("// Generated by tools/mt_jump.py: t^(c * stride) mod phi(t), phi = characteristic polynomial of MT19937,\n")
        f.write(f"// bit j of word j/32 = coefficient of t^j.  stride = {stride}, c = 1..{count}.\n")
        f.write(f"static const unsigned int kMtJumpStride = {stride}u;\n")
        f.write(f"static const unsigned int kMtJumpPoly[{count}][{words}] = {{\n")
        for c in range(1, count + 1):
            g = jump_polynomial(c * stride, phi)
            ws = [(g >> (32 * w)) & 0xFFFFFFFF for w in range(words)]
            f.write("  {" + ",".join(f"0x{w:08x}u" for w in ws) + "},\n")
        f.write("};\n")


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--header", default="")
    ap.add_argument("--stride", type=int, default=131072)
    ap.add_argument("--count", type=int, default=15)
    a = ap.parse_args()
    if a.check or not a.header:
        check()
    if a.header:
        write_header(a.header, a.stride, a.count)
