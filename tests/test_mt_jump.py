"""MT19937 jump-ahead identity (tools/mt_jump.py): groundwork for splitting the sampler's serial
draw generation over several CTAs.  CPU only."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import mt_jump  # noqa: E402


def test_word_sequence_is_mt19937():
    from oracle import sampler_oracle as so
    x = mt_jump.word_sequence(5489, 624 + 3000)
    assert np.array_equal(so._temper(x[624:]), so.mt19937_draws(5489, 3000))


@pytest.fixture(scope="module")
def phi():
    return mt_jump.characteristic_polynomial()


def test_characteristic_polynomial_is_the_known_one(phi):
    assert phi.bit_length() - 1 == 19937
    assert phi.bit_count() == 135                 # the published MT19937 characteristic polynomial has 135 terms
    assert phi & 1                                # nonsingular transition: constant term present


@pytest.mark.parametrize("J,seed", [(624, 0), (131072, 5489), (160146 * 3 + 7, 1 << 33), (1281166, 42)])
def test_jumped_window_equals_sequential_generation(phi, J, seed):
    g = mt_jump.jump_polynomial(J, phi)
    x = mt_jump.word_sequence(seed & 0xFFFFFFFF, J + 1400)
    got = mt_jump.jumped_window(x[: mt_jump.DEG + 1 + 624], g)
    assert np.array_equal(got, x[J + 1: J + 625])
    # the jumped window continues the recurrence exactly: 300 more words, and their tempered values are the draws
    ring = list(got.astype(np.uint64))
    for k in range(300):
        y = (int(ring[k]) & 0x80000000) | (int(ring[k + 1]) & 0x7FFFFFFF)
        ring.append(int(ring[k + 397]) ^ (y >> 1) ^ (0x9908B0DF if y & 1 else 0))
    assert np.array_equal(np.asarray(ring[624:], np.uint32), x[J + 625: J + 925])
