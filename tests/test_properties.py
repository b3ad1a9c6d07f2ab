"""Property tests (hypothesis) of host-side logic and of the oracle.  CPU only."""
import numpy as np
import torch
from hypothesis import assume, given, settings, strategies as st
from torch.utils.data.distributed import DistributedSampler

import b200ddp
from b200ddp.engine import padded_numel, split_buckets
from oracle import sampler_oracle as so


class _DS:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


@settings(max_examples=60, deadline=None, derandomize=True)
@given(st.lists(st.integers(0, 50_000), min_size=0, max_size=40), st.sampled_from([2, 4]),
       st.integers(4_096, 400_000), st.integers(200_000, 2_000_000))
def test_split_buckets_partitions_the_list(numels, esize, cap, staging):
    assume(all(padded_numel(n) * esize <= staging for n in numels))     # a tensor beyond the staging area is an error
    out = split_buckets(numels, esize, cap, staging)
    # consecutive, complete, in order
    assert [lo for lo, _, _, _ in out] == [0] + [hi for _, hi, _, _ in out][:-1] if out else numels == []
    if out:
        assert out[-1][1] == len(numels)
    for lo, hi, off, elems in out:
        assert hi > lo and off % 256 == 0
        assert elems == sum(padded_numel(n) for n in numels[lo:hi])
        assert off + elems * esize <= staging                       # every bucket fits the staging area
        # a bucket only exceeds the cap when it holds a single non-empty tensor (larger than the cap by itself)
        assert elems * esize <= min(cap, staging) or sum(1 for n in numels[lo:hi] if n > 0) == 1


@settings(max_examples=80, deadline=None, derandomize=True)
@given(st.integers(0, 3000), st.integers(1, 8), st.booleans(), st.booleans(), st.integers(0, 2 ** 40), st.integers(0, 5),
       st.data())
def test_sampler_oracle_equals_torch(n, world, shuffle, drop_last, seed, epoch, data):
    rank = data.draw(st.integers(0, world - 1))
    ref = DistributedSampler(_DS(n), num_replicas=world, rank=rank, shuffle=shuffle, seed=seed, drop_last=drop_last)
    ref.set_epoch(epoch)
    want = np.asarray(list(ref), dtype=np.int64)
    got = so.c_distributed_sampler_indices(n, world, rank, seed, epoch, shuffle, drop_last)
    assert got.shape == want.shape and np.array_equal(got, want)
    assert b200ddp.sampler_sizes(n, world, drop_last) == (ref.num_samples, ref.total_size)


@settings(max_examples=30, deadline=None, derandomize=True)
@given(st.integers(2, 8), st.integers(64, 400), st.integers(0, 2 ** 31))
def test_rank_order_sum_is_identical_on_every_rank_and_close_to_any_order(world, n, seed):
    """What the P2P kernels guarantee: a fixed (rank) summation order gives every rank the same bits,
    and differs from any other order by no more than the stated fp32 bound (inf-norm rel 1e-5)."""
    rng = np.random.default_rng(seed)
    g = [(rng.standard_normal(n) * 1e-2).astype(np.float32) for _ in range(world)]
    fixed = g[0].copy()
    for a in g[1:]:
        fixed = fixed + a
    perm = rng.permutation(world)
    other = g[perm[0]].copy()
    for i in perm[1:]:
        other = other + g[i]
    scale = max(np.max(np.abs(fixed)), 1e-30)
    assert np.max(np.abs(fixed - other)) <= 1e-5 * scale
    if world == 2:
        assert np.array_equal(fixed, other)                         # one addition: commutative, hence bit-exact
