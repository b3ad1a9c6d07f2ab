"""Real multi-process runs: one process per GPU, NCCL rendezvous, CUDA-IPC peer
memory - the launch model of codes/task2/model-mp.py:146-148.  Skipped on a box
with fewer GPUs than the test needs (the single-GPU emulation in
test_gpu_kernels.py covers the same kernels there)."""
import os
import traceback

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _entry(rank, ws, port, fn, args, q):
    import sys
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    try:
        import b200ddp
        b200ddp.dist_utils.dist_init(ws, rank, "127.0.0.1", str(port))
        fn(rank, ws, *args)
        torch.cuda.synchronize()
        q.put((rank, "ok"))
    except Exception:
        q.put((rank, traceback.format_exc()))
    finally:
        try:
            import b200ddp
            b200ddp.reset_default_context()
            torch.distributed.destroy_process_group()
        except Exception:
            pass


def _run(ws, fn, *args, timeout=240):
    if _ngpu() < ws:
        pytest.skip(f"needs {ws} GPUs, found {_ngpu()}")
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = 29500 + (os.getpid() * 7 + _run.n) % 400
    _run.n += 1
    procs = [ctx.Process(target=_entry, args=(r, ws, port, fn, args, q)) for r in range(ws)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout)
        if p.is_alive():
            p.kill()
            pytest.fail("multi-GPU worker timed out")
    res = dict(q.get() for _ in range(ws))
    bad = {r: m for r, m in res.items() if m != "ok"}
    assert not bad, "\n".join(f"rank {r}:\n{m}" for r, m in bad.items())


_run.n = 0


# ------------------------------------------------------------------ workers (module level: must pickle)
def _labnet_golden(rank, ws, golden, algo, mode, dtype):
    import b200ddp
    from oracle import ref_ddp
    g = dict(np.load(golden))
    numels, steps = ref_ddp.LABNET_NUMELS, int(g["steps"])
    splits = np.cumsum(numels)[:-1]
    b200ddp.configure(bucket_dtype=dtype, algo=algo)
    if algo == "nvls" and not b200ddp.default_context().has_multicast:
        return                                 # no multicast mapping on this box: nothing to test
    model = torch.nn.ParameterList(
        [torch.nn.Parameter(torch.from_numpy(a.copy()).cuda()) for a in np.split(g["p0"] + (rank != 0), splits)])
    b200ddp.init_parameters(model)            # rank 0's parameters everywhere (codes/task2/dist_utils.py:33-37)
    opt = b200ddp.SGD(model.parameters(), lr=0.01, momentum=0.9)
    if mode == "fused":
        b200ddp.enable_fused_step(model, opt)
    for s in range(steps):
        opt.zero_grad()
        for p, a in zip(model, np.split(g[f"local_r{rank}_s{s}"], splits)):
            p.grad = torch.from_numpy(a.copy()).cuda()
        b200ddp.allreduce_average_gradients(model)
        opt.step()
        avg = np.concatenate([p.grad.cpu().numpy() for p in model])
        ref = g[f"avg_s{s}"]
        if dtype == "bf16":
            assert np.max(np.abs(avg - ref)) <= 2.0 ** -7 * np.max(np.abs(ref))
        elif ws <= 2:
            assert np.array_equal(avg, ref), f"step {s}: averaged gradient differs"
        else:
            assert np.max(np.abs(avg - ref)) <= 1e-5 * np.max(np.abs(ref))
    p = np.concatenate([q.detach().cpu().numpy() for q in model])
    m = np.concatenate([opt.state[q]["momentum_buffer"].cpu().numpy() for q in model])
    if dtype == "bf16":
        assert np.max(np.abs(p - g["p_final"])) <= 1e-3 * np.max(np.abs(g["p_final"]))
    elif ws <= 2:
        assert np.array_equal(p, g["p_final"]) and np.array_equal(m, g["m_final"])
    else:
        assert np.max(np.abs(p - g["p_final"])) <= 1e-5 * np.max(np.abs(g["p_final"]))
        assert np.max(np.abs(m - g["m_final"])) <= 1e-5 * np.max(np.abs(g["m_final"]))
    # every rank holds the same bits
    mine = torch.from_numpy(p).cuda()
    ref0 = mine.clone()
    torch.distributed.broadcast(ref0, 0)
    assert torch.equal(mine, ref0), "ranks diverged"


def _grid_golden(rank, ws, golden, algo, mode):
    import b200ddp
    from oracle import synth
    g = dict(np.load(golden))
    numels, hp, steps = synth.RAGGED_NUMELS, synth.SGD_VARIANTS["A"], int(g["steps"])
    b200ddp.configure(bucket_dtype="fp32", algo=algo)
    model = torch.nn.ParameterList(
        [torch.nn.Parameter(torch.from_numpy(synth.init_param(n, t)).cuda()) for t, n in enumerate(numels)])
    opt = b200ddp.SGD(model.parameters(), **hp)
    if mode == "fused":
        b200ddp.enable_fused_step(model, opt)
    for s in range(steps):
        opt.zero_grad()
        for t, p in enumerate(model):
            p.grad = torch.from_numpy(synth.grid_grad(p.numel(), rank, s, t)).cuda()
        b200ddp.allreduce_average_gradients(model)
        opt.step()
    assert np.array_equal(np.concatenate([p.grad.cpu().numpy() for p in model]), g["avg"])
    assert np.array_equal(np.concatenate([p.detach().cpu().numpy() for p in model]), g["p"])
    assert np.array_equal(np.concatenate([opt.state[p]["momentum_buffer"].cpu().numpy() for p in model]), g["m"])


def _large_buckets(rank, ws):
    """ResNet-18-sized list through every algorithm: all must agree with NCCL's
    result within the fp32 bound and be identical across ranks; forces the
    multi-bucket path by shrinking the cap."""
    import torchvision
    import b200ddp
    from b200ddp.engine import GradSync
    torch.manual_seed(0)
    params = [torch.nn.Parameter(torch.zeros_like(p).cuda()) for p in torchvision.models.resnet18(num_classes=10).parameters()]
    gen = torch.Generator(device="cuda").manual_seed(100 + rank)
    local = [torch.randn(p.shape, device="cuda", generator=gen) * 1e-2 for p in params]
    results = {}
    cases = [("nccl", None), ("twoshot", None), ("oneshot", None), ("twoshot", 4 << 20), ("auto", 1 << 20)]
    if b200ddp.default_context().has_multicast:
        cases += [("nvls", None), ("nvls", 4 << 20)]
    for algo, cap in cases:
        eng = GradSync(params, b200ddp.default_context(), algo=algo, bucket_cap_bytes=cap)
        for p, gr in zip(params, local):
            p.grad = gr.clone()
        eng.sync()
        results[(algo, cap)] = torch.cat([p.grad.flatten() for p in params])
        eng.close()
    ref = results[("nccl", None)]
    for k, v in results.items():
        assert (v - ref).abs().max() <= 1e-5 * ref.abs().max(), k
        if k[0] not in ("nccl", "nvls") or (k[0] == "nvls" and ws == 2):
            assert torch.equal(v, results[("twoshot", None)]), k     # every P2P variant sums in rank order
        mine = v.clone()
        torch.distributed.broadcast(mine, 0)
        assert torch.equal(mine, v), f"{k}: ranks hold different sums"


def _ddp_hook(rank, ws):
    import copy
    import b200ddp
    from oracle import ref_ddp
    from torch.nn.parallel import DistributedDataParallel as DDP
    torch.manual_seed(0)
    # linear layers only: cuDNN's convolution backward is not run-to-run deterministic, and this test
    # compares two separate backward passes bit for bit
    base = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(784, 120), torch.nn.ReLU(),
                               torch.nn.Linear(120, 10)).cuda()
    m_hook, m_ref = copy.deepcopy(base), copy.deepcopy(base)
    d_hook = DDP(m_hook, device_ids=[torch.cuda.current_device()])
    d_hook.register_comm_hook(b200ddp.HookState(), b200ddp.comm_hook)
    d_ref = DDP(m_ref, device_ids=[torch.cuda.current_device()])
    loss_fn = torch.nn.CrossEntropyLoss()
    for s in range(3):
        x, y = ref_ddp.synthetic_batch(rank, s, 32, (1, 28, 28), 10)
        x, y = x.cuda(), y.cuda()
        # local gradients (no communication), to form the reference's own average by hand
        with d_ref.no_sync():
            d_ref.zero_grad()
            loss_fn(d_ref(x), y).backward()
        want = []
        for p in m_ref.parameters():
            parts = [torch.empty_like(p.grad) for _ in range(ws)]
            torch.distributed.all_gather(parts, p.grad)
            acc = parts[0].clone()
            for t in parts[1:]:
                acc += t
            want.append(acc / ws)            # codes/task2/dist_utils.py:41-42, summed in rank order
        d_hook.zero_grad()
        loss_fn(d_hook(x), y).backward()
        for a, w in zip(m_hook.parameters(), want):
            if ws <= 2:
                assert torch.equal(a.grad, w)
            else:
                assert (a.grad - w).abs().max() <= 1e-5 * w.abs().max()


def _graph_step(rank, ws, algo):
    """GraphedTrainStep across ranks: the captured step contains the P2P kernels, whose
    barrier generations live on the device - replays must stay in step with each other
    and reproduce the eager loop bit for bit."""
    import b200ddp
    from oracle import ref_ddp
    b200ddp.configure(algo=algo)
    outs = []
    for graphed in (False, True):
        torch.manual_seed(5)
        model = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(784, 300), torch.nn.ReLU(),
                                    torch.nn.Linear(300, 10)).cuda()
        b200ddp.init_parameters(model)
        opt = b200ddp.SGD(model.parameters(), lr=0.01, momentum=0.9)
        b200ddp.enable_fused_step(model, opt)
        loss_fn = torch.nn.CrossEntropyLoss()
        step = b200ddp.GraphedTrainStep(model, loss_fn, opt, eager_steps=2) if graphed else None
        for s in range(8):
            x, y = ref_ddp.synthetic_batch(rank, s, 32, (1, 28, 28), 10)
            x, y = x.cuda(), y.cuda()
            if graphed:
                step(x, y)
            else:
                loss = loss_fn(model(x), y)
                opt.zero_grad()
                loss.backward()
                b200ddp.allreduce_average_gradients(model)
                opt.step()
        torch.cuda.synchronize()
        if graphed:
            assert step.capture_error is None and step.graph is not None, step.capture_error
        outs.append(torch.cat([p.detach().flatten() for p in model.parameters()]))
    assert torch.equal(outs[0], outs[1]), "graph replay diverged from the eager loop"
    mine = outs[1].clone()
    torch.distributed.broadcast(mine, 0)
    assert torch.equal(mine, outs[1]), "ranks diverged"


def _next_rows(rank, ws):
    """flat initial-weight broadcast (with and without buffers), the
    all-gather aggregation variant, the CUDA-event comm timer, both sampler modes."""
    import b200ddp
    du = b200ddp.dist_utils

    def build(seed):
        torch.manual_seed(seed)
        return torch.nn.Sequential(torch.nn.Conv2d(1, 8, 3), torch.nn.BatchNorm2d(8), torch.nn.ReLU(),
                                   torch.nn.Flatten(), torch.nn.Linear(8 * 26 * 26, 10)).cuda()

    def same_everywhere(t):
        ref = t.detach().clone()
        torch.distributed.broadcast(ref, 0)
        return torch.equal(ref, t.detach())

    # (f1) init_parameters: parameters only, exactly like the reference ...
    m = build(10 + rank)
    with torch.no_grad():
        m[1].running_mean.fill_(float(rank))
    du.init_parameters(m)
    assert all(same_everywhere(p) for p in m.parameters())
    assert float(m[1].running_mean[0]) == float(rank)          # buffers untouched (codes/task2/dist_utils.py:35-37)
    # ... and the fixed variant that also sends buffers (fp32 through the flat path, int64 through torch)
    m = build(20 + rank)
    with torch.no_grad():
        m[1].running_mean.fill_(float(rank))
        m[1].num_batches_tracked.fill_(rank)
    du.init_parameters(m, buffers=True)
    assert all(same_everywhere(t) for t in list(m.parameters()) + list(m.buffers()))
    assert float(m[1].running_mean[0]) == 0.0 and int(m[1].num_batches_tracked) == 0

    # (f3) all-gather-then-mean aggregation == the mean over ranks, and the comm timer
    gen = torch.Generator(device="cuda").manual_seed(50 + rank)
    for p in m.parameters():
        p.grad = torch.randn(p.shape, device="cuda", generator=gen) * 1e-2
    want = []
    for p in m.parameters():
        parts = [torch.empty_like(p.grad) for _ in range(ws)]
        torch.distributed.all_gather(parts, p.grad)
        acc = parts[0].clone()
        for t in parts[1:]:
            acc += t
        want.append(acc / ws)
    timer = b200ddp.CommTimer()
    with timer:
        du.allgather_average_gradients(m)
    assert timer.total_seconds() > 0
    for p, w in zip(m.parameters(), want):
        if ws <= 2:
            assert torch.equal(p.grad, w)
        else:
            assert (p.grad - w).abs().max() <= 1e-5 * w.abs().max()

    # (f2) sampler modes of the lab: random partition (shared seed, disjoint shards) and random sampling
    class DS:
        def __len__(self):
            return 1000
    part = list(b200ddp.MySampler(DS(), ws, rank, shuffle=True, seed=0))
    gathered = [None] * ws
    torch.distributed.all_gather_object(gathered, part)
    flat = [i for shard in gathered for i in shard]
    assert sorted(set(flat)) == list(range(1000)) and len(flat) == -(-1000 // ws) * ws
    samp = list(b200ddp.MySampler(DS(), ws, rank, shuffle=True, seed=rank, mode="sampling"))   # codes/task3/model.py:111
    want_s = torch.randperm(1000, generator=torch.Generator().manual_seed(rank))[: -(-1000 // ws)].tolist()
    assert samp == want_s


# ------------------------------------------------------------------ tests
def test_next_rows_two_ranks():
    _run(2, _next_rows)


@pytest.mark.parametrize("algo", ["oneshot", "twoshot", "nccl"])
def test_graphed_step_two_ranks(algo):
    _run(2, _graph_step, algo)


def test_graphed_step_eight_ranks():
    _run(8, _graph_step, "auto")


@pytest.mark.parametrize("algo,mode,dtype", [("nvls", "strict", "fp32"), ("nvls", "fused", "bf16"),
                                             ("nccl", "strict", "fp32"), ("oneshot", "strict", "fp32"),
                                             ("twoshot", "strict", "fp32"), ("oneshot", "fused", "fp32"),
                                             ("twoshot", "fused", "fp32"), ("auto", "fused", "bf16")])
def test_labnet_golden_two_ranks(golden_dir, algo, mode, dtype):
    _run(2, _labnet_golden, os.path.join(golden_dir, "ddp_labnet_ws2.npz"), algo, mode, dtype)


@pytest.mark.parametrize("algo,mode", [("nccl", "strict"), ("oneshot", "fused"), ("twoshot", "strict")])
def test_labnet_golden_four_ranks(golden_dir, algo, mode):
    _run(4, _labnet_golden, os.path.join(golden_dir, "ddp_labnet_ws4.npz"), algo, mode, "fp32")


@pytest.mark.parametrize("ws", [2, 4, 8])
@pytest.mark.parametrize("algo,mode", [("oneshot", "fused"), ("twoshot", "strict"), ("nccl", "strict")])
def test_grid_golden_bit_exact(golden_dir, ws, algo, mode):
    _run(ws, _grid_golden, os.path.join(golden_dir, f"ddp_grid_A_ws{ws}.npz"), algo, mode)


@pytest.mark.parametrize("ws", [2, 8])
def test_resnet18_sized_buckets_all_algorithms(ws):
    _run(ws, _large_buckets)


def test_ddp_comm_hook_two_ranks():
    _run(2, _ddp_hook)


# ------------------------------------------------------------------ round 2: overlap, sub-groups, odd worlds, lab optimizers
def _overlap_equals_plain(rank, ws, graphed):
    """Buckets that leave from inside backward (reverse-order buckets, side stream) must give the bits
    the plain after-backward sync gives: same sums, same update, eager and replayed from a CUDA graph."""
    import b200ddp
    from oracle import ref_ddp
    outs = []
    for overlap in (False, True):
        b200ddp.configure(algo="auto", overlap=overlap, bucket_cap_bytes=(1 << 20) if overlap else 0)
        torch.manual_seed(5)
        # linear layers: their backward is run-to-run deterministic, and four of them give several buckets
        model = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(784, 512), torch.nn.ReLU(),
                                    torch.nn.Linear(512, 512), torch.nn.ReLU(), torch.nn.Linear(512, 256),
                                    torch.nn.ReLU(), torch.nn.Linear(256, 10)).cuda()
        b200ddp.init_parameters(model)
        opt = b200ddp.SGD(model.parameters(), lr=0.01, momentum=0.9)
        b200ddp.enable_fused_step(model, opt)
        loss_fn = torch.nn.CrossEntropyLoss()
        step = b200ddp.GraphedTrainStep(model, loss_fn, opt, eager_steps=2) if graphed else None
        for s in range(7):
            x, y = ref_ddp.synthetic_batch(rank, s, 32, (1, 28, 28), 10)
            x, y = x.cuda(), y.cuda()
            if graphed:
                step(x, y)
            else:
                loss = loss_fn(model(x), y)
                opt.zero_grad()
                loss.backward()
                b200ddp.allreduce_average_gradients(model)
                opt.step()
        torch.cuda.synchronize()
        eng = b200ddp.dist_utils.engine_for(model)
        if overlap:
            assert eng.overlap and len(eng.buckets) >= 3 and eng.buckets[0].hi == len(eng.params)
        if graphed:
            assert step.capture_error is None and step.graph is not None, step.capture_error
        outs.append(torch.cat([p.detach().flatten() for p in model.parameters()] +
                              [p.grad.flatten() for p in model.parameters()]))
    b200ddp.configure(overlap=False, bucket_cap_bytes=0)
    assert torch.equal(outs[0], outs[1]), "overlapped buckets changed the result"
    mine = outs[1].clone()
    torch.distributed.broadcast(mine, 0)
    assert torch.equal(mine, outs[1]), "ranks diverged"


def _overlap_strict_and_stale(rank, ws):
    """Strict mode with overlap (p.grad <- mean after the call), and the fallback when gradients are
    accumulated twice before the sync (what left early is stale and must not be used)."""
    import b200ddp
    b200ddp.configure(algo="auto", overlap=True, bucket_cap_bytes=1 << 20)
    torch.manual_seed(1)
    model = torch.nn.Sequential(torch.nn.Linear(600, 600), torch.nn.ReLU(), torch.nn.Linear(600, 600),
                                torch.nn.ReLU(), torch.nn.Linear(600, 10)).cuda()
    b200ddp.init_parameters(model)
    gen = torch.Generator(device="cuda").manual_seed(9 + rank)
    for accumulate in (1, 2, 1):
        model.zero_grad(set_to_none=True)
        for _ in range(accumulate):
            x = torch.randn(16, 600, device="cuda", generator=gen)
            model(x).square().mean().backward()
        local = [p.grad.clone() for p in model.parameters()]
        b200ddp.allreduce_average_gradients(model)
        for p, g in zip(model.parameters(), local):
            ref = g.clone()
            torch.distributed.all_reduce(ref)
            ref /= ws
            assert (p.grad - ref).abs().max() <= 1e-5 * ref.abs().max().clamp_min(1e-30)
            if ws == 2:
                assert torch.equal(p.grad, ref)
    b200ddp.configure(overlap=False, bucket_cap_bytes=0)


def _subgroups(rank, ws):
    """replicas of a pipeline stage sync over their own sub-group.  Ranks are split
    into `ws // 2` replicas x 2 stages (stage = rank % 2); each stage's gradients are averaged over the ranks
    holding that stage, on contexts connected over that sub-group."""
    import b200ddp
    from b200ddp.engine import GradSync
    stages = 2
    replicas = ws // stages
    groups = [torch.distributed.new_group([r for r in range(ws) if r % stages == s]) for s in range(stages)]
    stage = rank % stages
    grp = groups[stage]
    ctx = b200ddp.Context(rank // stages, replicas, torch.cuda.current_device(), staging_bytes=16 << 20)
    ctx.connect(group=grp)
    numels = (150, 6, 2400, 16) if stage == 0 else (48000, 120, 1200, 10)      # the task4 split, codes/task4/model.py:18-46
    params = [torch.nn.Parameter(torch.full((n,), float(stage), device="cuda")) for n in numels]
    gen = torch.Generator(device="cuda").manual_seed(30 + rank)
    for algo in ("twoshot", "oneshot", "nccl", "auto"):
        eng = GradSync(params, ctx, algo=algo)
        local = [torch.randn(n, device="cuda", generator=gen) * 1e-2 for n in numels]
        for p, g in zip(params, local):
            p.grad = g.clone()
        eng.sync()
        for p, g in zip(params, local):
            parts = [torch.empty_like(g) for _ in range(replicas)]
            torch.distributed.all_gather(parts, g, group=grp)
            acc = parts[0].clone()
            for t in parts[1:]:
                acc += t
            acc /= replicas
            if algo != "nccl":
                assert torch.equal(p.grad, acc), (algo, "sub-group mean differs")
            else:
                assert (p.grad - acc).abs().max() <= 1e-5 * acc.abs().max()
        eng.close()
    # the flat broadcast inside the sub-group: replica 0 of each stage is the source
    src = [torch.full((n,), 100.0 * stage + rank, device="cuda") for n in numels]
    b200ddp.broadcast_from_rank0(src, ctx)
    assert all(float(t[0]) == 100.0 * stage + stage for t in src)          # global rank of replica 0 of this stage = stage
    torch.cuda.synchronize()
    torch.distributed.barrier()
    ctx.close()


def _odd_world(rank, ws):
    """Three ranks: the sum in rank order, then the division as torch rounds it on CUDA - bit for bit
    against torch's own arithmetic on gathered gradients (codes/task2/dist_utils.py:41-42 run on the GPU)."""
    import b200ddp
    from b200ddp.engine import GradSync
    numels = (1000, 4099, 33, 70_000)
    params = [torch.nn.Parameter(torch.zeros(n, device="cuda")) for n in numels]
    gen = torch.Generator(device="cuda").manual_seed(70 + rank)
    for algo in ("twoshot", "oneshot", "auto"):
        local = [torch.randn(n, device="cuda", generator=gen) * 1e-2 for n in numels]
        want = []
        for g in local:
            parts = [torch.empty_like(g) for _ in range(ws)]
            torch.distributed.all_gather(parts, g)
            acc = parts[0].clone()
            for t in parts[1:]:
                acc += t
            acc /= ws
            want.append(acc)
        eng = GradSync(params, b200ddp.default_context(), algo=algo)
        for p, g in zip(params, local):
            p.grad = g.clone()
        eng.sync()
        assert all(torch.equal(p.grad, w) for p, w in zip(params, want)), algo
        # fused with the update: the division happens inside K3
        with torch.no_grad():
            for p in params:
                p.fill_(0.5)
        for p, g in zip(params, local):
            p.grad = g.clone()
        moms = [torch.empty_like(p) for p in params]
        eng.sync_and_step(moms, dict(lr=0.1, momentum=0.9), first_step=True)
        for p, m, w in zip(params, moms, want):
            assert torch.equal(p.grad, w) and torch.equal(m, w)
            # p + (-lr) * g as ONE fused multiply-add (torch/optim/sgd.py:380), formed in fp64
            fma = (0.5 + w.double() * torch.tensor(-0.1, dtype=torch.float32).double()).float()
            assert torch.equal(p.detach(), fma)
        eng.close()


def _lab_optimizers_fused(rank, ws):
    """The lab's Adam / GD fused with the sync (pack -> reduce -> unpack + rule) against the reference's
    statements run by torch on the averaged gradients."""
    import b200ddp
    from oracle import ref_ddp
    for cls in ("gd", "adam"):
        torch.manual_seed(3)
        model = torch.nn.Sequential(torch.nn.Linear(300, 200), torch.nn.ReLU(), torch.nn.Linear(200, 10)).cuda()
        b200ddp.configure(algo="auto")
        b200ddp.init_parameters(model)
        params = list(model.parameters())
        opt = b200ddp.GdOptimizer(params, lr=0.05) if cls == "gd" else b200ddp.AdamOptimizer(params, lr=0.002)
        b200ddp.enable_fused_step(model, opt)
        pt = [p.detach().clone() for p in params]
        mt, vt = [torch.zeros_like(p) for p in params], [torch.zeros_like(p) for p in params]
        gen = torch.Generator(device="cuda").manual_seed(11 + rank)
        for s in range(3):
            for p in params:
                p.grad = torch.randn(p.shape, device="cuda", generator=gen) * 1e-2
            avg = []
            for p in params:
                parts = [torch.empty_like(p.grad) for _ in range(ws)]
                torch.distributed.all_gather(parts, p.grad)
                acc = parts[0].clone()
                for t in parts[1:]:
                    acc += t
                avg.append(acc / ws)
            b200ddp.allreduce_average_gradients(model)
            opt.step()
            for i in range(len(params)):
                if cls == "gd":
                    pt[i] = ref_ddp.gd_step_torch(pt[i], avg[i], lr=0.05)
                else:
                    pt[i], mt[i], vt[i] = ref_ddp.adam_step_torch(pt[i], avg[i], mt[i], vt[i], lr=0.002)
            for p, a, w in zip(params, avg, pt):
                assert torch.equal(p.grad, a) if ws == 2 else (p.grad - a).abs().max() <= 1e-5 * a.abs().max()
                if ws == 2:
                    assert torch.equal(p.detach(), w), cls
                else:
                    assert (p.detach() - w).abs().max() <= 1e-5 * w.abs().max()


@pytest.mark.parametrize("graphed", [False, True])
def test_overlap_equals_plain_two_ranks(graphed):
    _run(2, _overlap_equals_plain, graphed)


def test_overlap_equals_plain_eight_ranks_graphed():
    _run(8, _overlap_equals_plain, True)


def test_overlap_strict_and_stale_two_ranks():
    _run(2, _overlap_strict_and_stale)


def test_subgroup_contexts_four_ranks():
    _run(4, _subgroups)


def test_subgroup_contexts_eight_ranks():
    _run(8, _subgroups)


def test_three_ranks_sum_then_divide_bit_exact():
    _run(3, _odd_world)


def test_lab_optimizers_fused_two_ranks():
    _run(2, _lab_optimizers_fused)


@pytest.mark.parametrize("ws", [4, 8])
def test_ddp_comm_hook_more_ranks(ws):
    _run(ws, _ddp_hook)


def test_next_rows_four_ranks():
    _run(4, _next_rows)


def _task4_dp(rank, ws, microbatches):
    """R replicas x 2 stages (BASELINE.json configs[3]) against plain data parallelism of the WHOLE lab
    CNN over R ranks, computed locally by every rank from all replicas' batches: same split
    (codes/task4/model.py:18-46), plain SGD lr=0.01 (:126), gradient = mean over replicas
    (codes/task4/dist_utils.py:40-46)."""
    import b200ddp
    from oracle import ref_ddp
    R = ws // 2
    rep = b200ddp.TwoStageReplica(lr=0.01, microbatches=microbatches, seed=4)
    torch.manual_seed(4)
    full = ref_ddp.LabNet().cuda()
    loss_fn = torch.nn.CrossEntropyLoss()
    for s in range(4):
        batches = [ref_ddp.synthetic_batch(r, s, 32, (1, 28, 28), 10) for r in range(R)]
        x, y = batches[rep.replica]
        loss = rep.step(inputs=x.cuda() if rep.stage == 0 else None, labels=y.cuda() if rep.stage == 1 else None)
        # expected: mean over replicas of the full model's gradients, then p <- p - lr * g
        grads = []
        for bx, by in batches:
            full.zero_grad(set_to_none=True)
            loss_fn(full(bx.cuda()), by.cuda()).backward()
            grads.append([p.grad.clone() for p in full.parameters()])
        with torch.no_grad():
            for i, p in enumerate(full.parameters()):
                acc = grads[0][i].clone()
                for g in grads[1:]:
                    acc += g[i]
                p.add_(acc / R, alpha=-0.01)
        if rep.stage == 1:
            assert loss is not None and torch.isfinite(loss)
    want = list(full.parameters())[:4] if rep.stage == 0 else list(full.parameters())[4:]
    for p, w in zip(rep.params, want):
        assert p.shape == w.shape
        assert (p.detach() - w.detach()).abs().max() <= 1e-5 * w.detach().abs().max(), "stage diverged from full-model data parallelism"
    # replicas of a stage hold identical bits
    flat = torch.cat([p.detach().flatten() for p in rep.params])
    ref0 = flat.clone()
    torch.distributed.broadcast(ref0, rep.stage, group=rep.group)          # global rank of replica 0 of this stage
    assert torch.equal(flat, ref0), "replicas diverged"
    torch.cuda.synchronize()
    torch.distributed.barrier()
    rep.close()


@pytest.mark.parametrize("ws,microbatches", [(2, 1), (4, 1), (4, 4), (8, 2)])
def test_task4_two_stage_replicated(ws, microbatches):
    _run(ws, _task4_dp, microbatches)
