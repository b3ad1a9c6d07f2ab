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
