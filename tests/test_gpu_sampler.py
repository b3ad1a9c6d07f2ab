"""K4: the on-device DistributedSampler shard against the reference-generated
fixtures and the CPU oracle.  Integer work: everything is bit-exact."""
import hashlib
import os

import numpy as np
import pytest
import torch

import b200ddp
from oracle import sampler_oracle as so

pytestmark = pytest.mark.gpu


def _shard(ctx, n, N, r, seed=0, epoch=0, shuffle=True, drop_last=False):
    return ctx.sampler_shard(n, N, r, seed, epoch, shuffle, drop_last).cpu().numpy()


def test_small_golden_all_cases(ctx1, golden_dir):
    g = dict(np.load(os.path.join(golden_dir, "sampler_small.npz")))
    for key, ref in g.items():
        f = {kv[0]: int(kv[1:]) for kv in key.split("_")}
        got = _shard(ctx1, f["n"], f["N"], f["r"], 3, f["e"], bool(f["s"]), bool(f["d"]))
        assert got.dtype == np.int64 and np.array_equal(got, ref.astype(np.int64)), key


def test_reference_configuration(ctx1, golden_dir):
    # DistributedSampler(train_set, num_replicas=2, rank) on MNIST: codes/task2/model.py:124
    g = dict(np.load(os.path.join(golden_dir, "sampler_mnist_ws2.npz")))
    for r in range(2):
        assert np.array_equal(_shard(ctx1, 60000, 2, r), g[f"r{r}"].astype(np.int64))


def test_big_golden_digests(ctx1, golden_dir):
    g = dict(np.load(os.path.join(golden_dir, "sampler_big.npz")))
    for key in sorted(k[:-4] for k in g if k.endswith("_sha")):
        parts = key.split("_")
        n, N = int(parts[0][1:]), int(parts[1][1:])
        seed, e, d, r = int(parts[2][4:]), int(parts[3][1:]), int(parts[4][1:]), int(parts[5][1:])
        got = _shard(ctx1, n, N, r, seed, e, True, bool(d))
        assert len(got) == int(g[key + "_len"])
        assert hashlib.sha256(got.tobytes()).digest() == bytes(g[key + "_sha"]), key


@pytest.mark.parametrize("n", [1, 2, 3, 226, 227, 228, 623, 624, 625, 1023, 1024, 1025, 2049, 4097, 65535, 65536,
                               65537, 100003, 303104, 1281167])
def test_full_permutation_matches_oracle(ctx1, n):
    # world 1 returns the whole permutation; sizes straddle the MT block (227/624) and grid boundaries
    for seed in (0, 12345):
        assert np.array_equal(_shard(ctx1, n, 1, 0, seed), so.c_randperm(n, seed)), (n, seed)


def test_is_a_permutation_and_shards_partition_it(ctx1):
    n, N = 1281167, 8          # ImageNet-sized, 8 ranks (BASELINE.json config 3)
    shards = [_shard(ctx1, n, N, r, 0, 5) for r in range(N)]
    assert all(len(s) == 160146 for s in shards)
    allidx = np.stack(shards, 1).reshape(-1)            # position t = rank + N * k
    assert np.array_equal(np.sort(allidx[:n]), np.arange(n))
    assert np.array_equal(allidx[n:], allidx[: N * 160146 - n])   # pad repeats the head


def test_seed_uses_low_32_bits_like_torch(ctx1):
    big = (7 << 32) + 99
    ref = torch.randperm(5000, generator=torch.Generator().manual_seed(big)).numpy()
    assert np.array_equal(_shard(ctx1, 5000, 1, 0, big), ref)
    assert np.array_equal(_shard(ctx1, 5000, 1, 0, 99), ref)


def test_epoch_adds_to_seed(ctx1):
    assert np.array_equal(_shard(ctx1, 10007, 4, 1, seed=3, epoch=4), _shard(ctx1, 10007, 4, 1, seed=7, epoch=0))
    assert not np.array_equal(_shard(ctx1, 10007, 4, 1, seed=3, epoch=4), _shard(ctx1, 10007, 4, 1, seed=3, epoch=0))


def test_scratch_grows_and_shrinks_between_calls(ctx1):
    for n in (100, 200000, 50, 60000, 200000):
        assert np.array_equal(_shard(ctx1, n, 2, 1, 1), so.c_distributed_sampler_indices(n, 2, 1, 1))


def test_repeated_calls_are_identical(ctx1):
    a = _shard(ctx1, 60000, 2, 0)
    for _ in range(5):
        assert np.array_equal(_shard(ctx1, 60000, 2, 0), a)


def test_bad_arguments(ctx1):
    with pytest.raises(b200ddp.B200DDPError, match="Invalid rank"):
        ctx1.sampler_shard(10, 2, 2, 0, 0, True, False)
    with pytest.raises(b200ddp.B200DDPError, match="32-bit randperm range"):
        ctx1.sampler_shard((1 << 32) // 20, 1, 0, 0, 0, True, False)


def test_public_sampler_classes_match_torch(ctx1):
    from torch.utils.data.distributed import DistributedSampler as TorchDS
    b200ddp.runtime._default = ctx1          # use the fixture's context as the process default
    try:
        class DS:
            def __len__(self):
                return 10007
        for N, r, dl in ((2, 1, False), (4, 3, True), (8, 0, False)):
            mine = b200ddp.DistributedSampler(DS(), num_replicas=N, rank=r, seed=5, drop_last=dl)
            ref = TorchDS(DS(), num_replicas=N, rank=r, seed=5, drop_last=dl)
            for epoch in (0, 2):
                mine.set_epoch(epoch)
                ref.set_epoch(epoch)
                assert list(mine) == list(ref) and len(mine) == len(ref)
        with pytest.raises(ValueError, match="Invalid rank"):
            b200ddp.DistributedSampler(DS(), num_replicas=2, rank=2)
        # the reference skeleton's constructor, both lab modes (codes/task3/sampler.py:5-14, codes/task3/model.py:111)
        part = b200ddp.MySampler(DS(), 4, 2, shuffle=True, seed=0)
        assert list(part) == list(TorchDS(DS(), num_replicas=4, rank=2, seed=0)) and len(part) == 2502
        samp = b200ddp.MySampler(DS(), 4, 2, shuffle=True, seed=2, mode="sampling")
        assert list(samp) == so.c_randperm(10007, 2)[:2502].tolist()
        loader = torch.utils.data.DataLoader(torch.arange(10007), batch_size=32, sampler=part)
        assert sum(len(b) for b in loader) == 2502
    finally:
        b200ddp.runtime._default = None


def test_cluster_and_grid_kernels_agree(ctx1):
    """n <= 65,535 runs in one thread-block cluster (state in distributed shared memory), larger n
    in the cooperative grid kernel; B200DDP_SAMPLER_NO_CLUSTER forces the latter.  Both must give
    torch's permutation."""
    for n in (1, 2, 8, 9, 1000, 8191, 8192, 8193, 50000, 60000, 65535):
        for seed in (0, 77):
            os.environ.pop("B200DDP_SAMPLER_NO_CLUSTER", None)
            a = _shard(ctx1, n, 1, 0, seed)
            os.environ["B200DDP_SAMPLER_NO_CLUSTER"] = "1"
            try:
                b = _shard(ctx1, n, 1, 0, seed)
            finally:
                os.environ.pop("B200DDP_SAMPLER_NO_CLUSTER", None)
            ref = so.c_randperm(n, seed)
            assert np.array_equal(a, ref), (n, seed, "cluster")
            assert np.array_equal(b, ref), (n, seed, "grid")
