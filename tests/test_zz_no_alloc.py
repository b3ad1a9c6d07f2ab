"""b200ddp_sampler_shard_static: the sampler shard for callers that must not
allocate (their own pool owns the device memory, or they cannot take
cudaMalloc's device-wide stall).  Called through the C ABI and compared with
the regular entry point, which has already been checked against torch, and
directly with the committed torch golden vector and the C restatement.

Integer output, so the comparison is exact - the tolerance every sampler test
in this suite uses.  This file sorts last on purpose: a `-x` run reaches it only
after everything else has run.
"""
import ctypes
import os
from ctypes import c_void_p

import numpy as np
import pytest
import torch

import b200ddp
from b200ddp import _lib

pytestmark = pytest.mark.gpu


def _static_call(out, n, world, rank, seed, epoch, shuffle=True, drop_last=False):
    """Straight through ctypes: no context, no library-side allocation."""
    return _lib.lib().b200ddp_sampler_shard_static(
        n, world, rank, seed, epoch, int(shuffle), int(drop_last), c_void_p(out.data_ptr()),
        c_void_p(torch.cuda.current_stream().cuda_stream))


def test_static_sampler_matches_regular_entry_point_twice(built_lib):
    torch.cuda.set_device(0)
    ctx = b200ddp.Context(0, 1, 0, staging_bytes=1 << 20)
    try:
        # caller-owned outputs, allocated up front
        out_a = torch.full((30000,), -7, dtype=torch.int64, device="cuda")
        out_b = torch.full((12500,), -7, dtype=torch.int64, device="cuda")
        # call 1: the reference's own configuration (MNIST length, 2 ranks); first call on this device does the set-up
        assert _static_call(out_a, 60000, 2, 1, 0, 0) == 0, _lib.last_error()
        want_a = ctx.sampler_shard(60000, 2, 1, 0, 0, True, False)
        assert torch.equal(out_a, want_a)
        # call 2, straight after, smaller n and another seed: nothing left in the shared scratch by call 1 may leak in
        free_before = torch.cuda.mem_get_info()[0]
        assert _static_call(out_b, 50000, 4, 3, 17, 2) == 0, _lib.last_error()
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before          # the call took no device memory
        want_b = ctx.sampler_shard(50000, 4, 3, 17, 2, True, False)
        assert torch.equal(out_b, want_b)
        # and the first configuration again gives the first answer again
        out_a.fill_(-7)
        assert _static_call(out_a, 60000, 2, 1, 0, 0) == 0
        assert torch.equal(out_a, want_a)
        # the largest accepted n, and the unshuffled path
        big = torch.empty(_lib.SAMPLER_STATIC_MAX_N, dtype=torch.int64, device="cuda")
        assert _static_call(big, _lib.SAMPLER_STATIC_MAX_N, 1, 0, 5, 0) == 0
        assert torch.equal(big, ctx.sampler_shard(_lib.SAMPLER_STATIC_MAX_N, 1, 0, 5, 0, True, False))
        assert _static_call(out_b, 50000, 4, 3, 0, 0, shuffle=False) == 0
        assert torch.equal(out_b, ctx.sampler_shard(50000, 4, 3, 0, 0, False, False))
    finally:
        torch.cuda.synchronize()
        ctx.close()


def test_static_sampler_matches_torch_golden_and_c_restatement(built_lib, golden_dir):
    """Not only equal to the library's other entry point: equal to what torch's DistributedSampler
    gave for the reference's own configuration (tests/golden/sampler_mnist_ws2.npz, made by
    oracle/make_golden.py), and to the plain-C restatement on a second geometry."""
    from oracle import sampler_oracle
    torch.cuda.set_device(0)
    g = np.load(os.path.join(golden_dir, "sampler_mnist_ws2.npz"))
    out = torch.empty(30000, dtype=torch.int64, device="cuda")
    for rank in (0, 1):
        assert _static_call(out, 60000, 2, rank, 0, 0) == 0, _lib.last_error()
        assert np.array_equal(out.cpu().numpy(), g[f"r{rank}"].astype(np.int64))
    for n, world, rank, seed, epoch, drop_last in ((50000, 8, 5, 3, 1, False), (10007, 3, 2, 0, 4, True), (65536, 4, 0, 9, 0, False)):
        want = sampler_oracle.c_distributed_sampler_indices(n, world, rank, seed=seed, epoch=epoch, drop_last=drop_last)
        got = torch.empty(len(want), dtype=torch.int64, device="cuda")
        assert _static_call(got, n, world, rank, seed, epoch, drop_last=drop_last) == 0, _lib.last_error()
        assert np.array_equal(got.cpu().numpy(), np.asarray(want, dtype=np.int64))


def test_static_sampler_refuses_what_does_not_fit(built_lib):
    torch.cuda.set_device(0)
    out = torch.empty(8, dtype=torch.int64, device="cuda")
    rc = _static_call(out, _lib.SAMPLER_STATIC_MAX_N + 1, 1, 0, 0, 0)
    assert rc == -5 and "exceeds" in _lib.last_error() and str(_lib.SAMPLER_STATIC_MAX_N) in _lib.last_error()
    assert _static_call(out, 10, 2, 2, 0, 0) == -1 and "Invalid rank" in _lib.last_error()
    with pytest.raises(b200ddp.B200DDPError, match="exceeds"):
        b200ddp.runtime.sampler_shard_static(out, 1281167, 8, 0)
    got = b200ddp.runtime.sampler_shard_static(out, 16, 2, 1, seed=3)      # Python wrapper, small case
    ctx = b200ddp.Context(0, 1, 0, staging_bytes=1 << 20)
    assert torch.equal(got, ctx.sampler_shard(16, 2, 1, 3, 0, True, False))
    ctx.close()
