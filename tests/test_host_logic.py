"""Host-side logic and CPU emulations of the device algorithms.  CPU only."""
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import b200ddp
from b200ddp.engine import padded_numel, split_buckets
from oracle import ref_ddp, sampler_oracle as so


# ------------------------------------------------------------------ bucket layout
def test_split_single_bucket_labnet():
    out = split_buckets(ref_ddp.LABNET_NUMELS, 4, None, 1 << 20)
    assert out == [(0, 8, 0, sum(padded_numel(n) for n in ref_ddp.LABNET_NUMELS))]


def test_split_respects_cap_and_alignment():
    numels = [1000, 7, 0, 5000, 33, 1]
    out = split_buckets(numels, 4, 8192, 1 << 20)
    assert [b[0] for b in out][0] == 0 and out[-1][1] == len(numels)
    for (lo, hi, off, elems), nxt in zip(out, out[1:] + [None]):
        assert off % 256 == 0
        assert elems == sum(padded_numel(n) for n in numels[lo:hi])
        if nxt:
            assert nxt[0] == hi
    # a tensor larger than the cap still gets a bucket of its own
    assert any(hi - lo == 1 and numels[lo] == 5000 for lo, hi, _, _ in out)


def test_split_reuses_staging_and_rejects_oversize():
    out = split_buckets([1000] * 10, 4, 8192, 20000)
    assert {b[2] for b in out} == {0, 8192}
    with pytest.raises(b200ddp.B200DDPError):
        split_buckets([10_000], 4, None, 4096)


def test_split_bf16_halves_the_bytes():
    a = split_buckets([4096] * 8, 4, 16384, 1 << 20)
    b = split_buckets([4096] * 8, 2, 16384, 1 << 20)
    assert len(a) == 8 and len(b) == 4


# ------------------------------------------------------------------ two-shot slice partition
@pytest.mark.parametrize("nvec,world", [(1, 2), (7, 8), (8, 8), (1000, 3), (13000, 8), (52000 // 4, 4)])
def test_twoshot_slices_cover_bucket_once(nvec, world):
    """csrc/p2p.cu twoshot_kernel: per = ceil(nvec / world); rank r owns [r*per, min((r+1)*per, nvec))."""
    per = (nvec + world - 1) // world
    seen = np.zeros(nvec, int)
    for r in range(world):
        lo, hi = per * r, min(per * r + per, nvec)
        if lo < hi:
            seen[lo:hi] += 1
    assert (seen == 1).all()


# ------------------------------------------------------------------ sampler: deterministic reservations
def _reservation_shuffle(n, seed, lag=0):
    """numpy emulation of csrc/sampler.cu: rounds of reserve (max priority =
    lowest pending iterate) / commit.  `lag` lets a fraction of iterates run one
    reserve phase ahead, as blocks do without a barrier after the commit."""
    draws = so.mt19937_draws(seed, max(n - 1, 0)).astype(np.int64)
    it = np.arange(n - 1, dtype=np.int64)
    tgt = it + draws % (n - it)
    perm = np.arange(n, dtype=np.int64)
    pending = np.ones(n - 1, bool)
    rounds = 0
    while pending.any():
        rounds += 1
        idx = it[pending]
        resv = np.full(n, n, np.int64)            # lower iterate = higher priority
        np.minimum.at(resv, idx, idx)
        np.minimum.at(resv, tgt[idx], idx)
        win = (resv[idx] == idx) & (resv[tgt[idx]] == idx)
        if lag:                                    # spurious failures must only delay, never corrupt
            win &= (np.arange(len(idx)) % lag) != 0
            if not win.any():
                win[0] = True                      # the lowest pending iterate always wins on hardware too
        w = idx[win]
        a, b = perm[w].copy(), perm[tgt[w]].copy()
        perm[w], perm[tgt[w]] = b, a
        pending[w] = False
    return perm, rounds


@pytest.mark.parametrize("n,seed", [(2, 0), (3, 1), (10, 0), (1000, 7), (60000, 0), (50000, 3)])
def test_reservation_shuffle_equals_sequential(n, seed):
    perm, rounds = _reservation_shuffle(n, seed)
    assert np.array_equal(perm, so.c_randperm(n, seed))
    assert rounds <= 6 * max(1, int(np.log2(n))) + 8          # O(log n) dependence depth


def test_reservation_shuffle_tolerates_spurious_failures():
    perm, _ = _reservation_shuffle(5000, 11, lag=3)
    assert np.array_equal(perm, so.c_randperm(5000, 11))


def test_mt_block_parallel_recurrence():
    """csrc/sampler.cu mt_targets: word k+624 from words k, k+1, k+397, 227 at a time."""
    seed, ndraw = 1234, 3000
    x = list(so.mt19937_init(seed).astype(np.uint64))
    for base in range(0, ndraw, 227):
        new = []
        for k in range(base, min(base + 227, ndraw)):
            y = (int(x[k]) & 0x80000000) | (int(x[k + 1]) & 0x7FFFFFFF)
            new.append(int(x[k + 397]) ^ (y >> 1) ^ (0x9908B0DF if y & 1 else 0))
        x.extend(new)                      # nothing computed in this step was read in this step
    got = so._temper(np.asarray(x[624:624 + ndraw], np.uint32))
    assert np.array_equal(got, so.mt19937_draws(seed, ndraw))


# ------------------------------------------------------------------ scale policy
@pytest.mark.parametrize("world", [1, 2, 3, 4, 5, 8])
def test_prescale_equals_postdivide_for_pow2(world):
    """Power-of-two worlds pre-scale in the pack; the result must equal the
    reference's sum-then-divide bit for bit (barring under/overflow)."""
    rng = np.random.default_rng(world)
    g = [(rng.standard_normal(4096) * 1e-2).astype(np.float32) for _ in range(world)]
    post = g[0].copy()
    for a in g[1:]:
        post = post + a
    post = post / np.float32(world)
    pre = g[0] * np.float32(1.0 / world)
    for a in g[1:]:
        pre = pre + a * np.float32(1.0 / world)
    if world & (world - 1) == 0:
        assert np.array_equal(pre, post)
    else:
        assert np.max(np.abs(pre - post)) <= 1e-6 * np.max(np.abs(post))


# ------------------------------------------------------------------ world_size-2 gloo run of the oracle
def _gloo_worker(rank, ws, port, golden, q):
    torch.set_num_threads(1)
    ref_ddp.dist_init(ws, rank, "127.0.0.1", str(port))
    g = dict(np.load(golden))
    splits = np.cumsum(ref_ddp.LABNET_NUMELS)[:-1]
    model = torch.nn.ParameterList([torch.nn.Parameter(torch.from_numpy(a.copy())) for a in np.split(g["p0"], splits)])
    opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9)
    ok = True
    for s in range(int(g["steps"])):
        for p, a in zip(model, np.split(g[f"local_r{rank}_s{s}"], splits)):
            p.grad = torch.from_numpy(a.copy())
        ref_ddp.allreduce_average_gradients(model)
        ok &= np.array_equal(np.concatenate([p.grad.numpy().ravel() for p in model]), g[f"avg_s{s}"])
        opt.step()
    ok &= np.array_equal(np.concatenate([p.detach().numpy().ravel() for p in model]), g["p_final"])
    # sampler shards of the two ranks partition the padded permutation
    mine = so.c_distributed_sampler_indices(1001, ws, rank, 5, 2)
    gathered = [None] * ws
    dist.all_gather_object(gathered, mine.tolist())
    allidx = np.asarray(gathered).T.reshape(-1)
    ok &= sorted(set(allidx.tolist())) == list(range(1001)) and len(allidx) == 1002
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_oracle_under_gloo_world2(golden_dir):
    import os
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = 29700 + os.getpid() % 200
    golden = os.path.join(golden_dir, "ddp_labnet_ws2.npz")
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, golden, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    res = dict(q.get() for _ in range(2))
    assert res == {0: True, 1: True}


# ------------------------------------------------------------------ short correctly-rounded division
def test_short_division_is_correctly_rounded():
    """csrc/device_utils.cuh div_rn: q = RN(x * rcp), r = fma(-d, q, x), q' = fma(r, rcp, q) must equal
    the IEEE quotient x / d for every integer divisor 2..64 (world sizes), as long as x is in the
    range the kernel takes the short path for.  fma is emulated through float64 (the product of two
    fp32 is exact there)."""
    def fma32(a, b, c):
        return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(np.float32)
    rng = np.random.default_rng(0)
    x = np.concatenate([
        (rng.standard_normal(400_000) * 10.0 ** rng.integers(-25, 25, 400_000)).astype(np.float32),
        np.arange(0x3F800000, 0x40000000, 37, dtype=np.uint32).view(np.float32),          # a sweep of mantissas
        np.float32([1e-30, 1.0000001e-30, 9.99e29, 3.0, -3.0, 1.0, 0.1, 7.0, 2.0 ** -20, 4096 * 2.0 ** -20])])
    x = x[(np.abs(x) > 1e-30) & (np.abs(x) < 1e30)]
    for d in range(2, 65):
        dd = np.float32(d)
        rcp = np.float32(1.0) / dd
        q = (x * rcp).astype(np.float32)
        q2 = fma32(fma32(np.full_like(x, -dd), q, x), np.full_like(x, rcp), q)
        assert np.array_equal(q2, (x / dd).astype(np.float32)), d


# ------------------------------------------------------------------ cross-GPU barrier protocol
@pytest.mark.parametrize("world,seed", [(2, 0), (4, 1), (8, 2), (3, 3)])
def test_flag_barrier_protocol_under_random_interleaving(world, seed):
    """Model of csrc/p2p.cu peer_barrier + device-side generations: flag[dst][block][src] holds the
    highest barrier generation block `block` of rank `src` has reached; block_gen[rank][block] is
    advanced by two per launch.  Ranks run the SAME sequence of launches (grid sizes vary), kernels
    of one rank run in stream order, everything else interleaves at random.  Must never deadlock,
    and a block may only pass barrier g after every peer's same block reached g."""
    rng = np.random.default_rng(seed)
    grids = [int(g) for g in rng.integers(1, 6, 12)]                 # 12 launches, 1..5 blocks each
    nblk = max(grids)
    flag = np.zeros((world, nblk, world), np.int64)
    gen = np.zeros((world, nblk), np.int64)
    reached = np.zeros((world, nblk), np.int64)                      # ground truth: barriers entered so far
    # per rank: index of the current launch and, per block of it, the phase 0..4
    launch = [0] * world
    phase = [[0] * grids[0] for _ in range(world)]
    base = [[int(gen[r][b]) for b in range(grids[0])] for r in range(world)]
    steps = 0
    while any(l < len(grids) for l in launch):
        steps += 1
        assert steps < 200000, "barrier protocol deadlocked"
        r = int(rng.integers(world))
        if launch[r] >= len(grids):
            continue
        b = int(rng.integers(grids[launch[r]]))
        ph, g0 = phase[r][b], base[r][b]
        if ph in (0, 2):                                             # signal barrier g0+1 / g0+2 to every peer
            g = g0 + 1 + ph // 2
            reached[r][b] = g
            flag[:, b, r] = g
            phase[r][b] += 1
        elif ph in (1, 3):                                           # wait for every peer's flag
            g = g0 + 1 + ph // 2
            if (flag[r, b, :] >= g).all():
                assert (reached[:, b] >= g).all(), "passed a barrier a peer had not reached"
                phase[r][b] += 1
                if phase[r][b] == 4:
                    gen[r][b] = g0 + 2
        if all(p == 4 for p in phase[r]):                            # kernel done: next launch of this rank's stream
            launch[r] += 1
            if launch[r] < len(grids):
                n = grids[launch[r]]
                phase[r] = [0] * n
                base[r] = [int(gen[r][bb]) for bb in range(n)]
    assert (gen == gen[0]).all()                                     # every rank ends on the same generations


# ------------------------------------------------------------------ sampler classes: argument handling (no GPU needed)
def test_sampler_constructors_match_torch_argument_handling():
    from torch.utils.data.distributed import DistributedSampler as TorchDS

    class DS:
        def __init__(self, n):
            self.n = n

        def __len__(self):
            return self.n

    for n, world, drop_last in ((60000, 2, False), (10007, 4, True), (5, 8, False), (3, 8, True), (0, 2, False)):
        mine = b200ddp.DistributedSampler(DS(n), num_replicas=world, rank=world - 1, drop_last=drop_last)
        ref = TorchDS(DS(n), num_replicas=world, rank=world - 1, drop_last=drop_last)
        assert (len(mine), mine.total_size, mine.epoch) == (len(ref), ref.total_size, 0)
        mine.set_epoch(7)
        assert mine.epoch == 7
    for bad in (-1, 2):
        with pytest.raises(ValueError, match="Invalid rank"):
            b200ddp.DistributedSampler(DS(10), num_replicas=2, rank=bad)
        with pytest.raises(ValueError, match="Invalid rank"):
            TorchDS(DS(10), num_replicas=2, rank=bad)
    with pytest.raises(RuntimeError):                       # no process group and no explicit replica count
        b200ddp.DistributedSampler(DS(10))
    # the reference skeleton's constructor (codes/task3/sampler.py:6-14): positional num_replicas, rank
    s = b200ddp.MySampler(DS(1001), 4, 1, shuffle=True, seed=3)
    assert (s.num_replicas, s.rank, s.shuffle, s.seed, s.epoch, len(s)) == (4, 1, True, 3, 0, 251)
    with pytest.raises(ValueError):
        b200ddp.MySampler(DS(10), 2, 0, mode="bogus")


@pytest.mark.parametrize("n", [2, 3, 4, 5, 6, 7, 8])
def test_oracle_division_rules(n):
    """The two roundings of `grad /= world_size`: identical for power-of-two worlds, at most one ulp
    apart otherwise (and really different somewhere, which is why the rule is an option)."""
    from oracle import ref_ddp
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(100_000) * 10.0 ** rng.integers(-20, 20, 100_000)).astype(np.float32)
    cpu, cuda = ref_ddp.divide_by_world(x, n, "cpu"), ref_ddp.divide_by_world(x, n, "cuda")
    assert np.array_equal(cpu, (torch.from_numpy(x) / n).numpy())          # ATen's CPU kernel is the true quotient
    if n & (n - 1) == 0:
        assert np.array_equal(cpu, cuda)
    else:
        ulp = np.abs(cpu.view(np.int32).astype(np.int64) - cuda.view(np.int32).astype(np.int64))
        assert ulp.max() == 1
