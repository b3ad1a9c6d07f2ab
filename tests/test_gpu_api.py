"""The reference-facing Python surface on one GPU: optimizer shim, dist_utils
names, DDP hook plumbing."""
import numpy as np
import pytest
import torch

import b200ddp
from oracle import ref_ddp, synth

pytestmark = pytest.mark.gpu


@pytest.fixture()
def default_ctx(ctx1):
    b200ddp.runtime._default = ctx1
    yield ctx1
    b200ddp.runtime._default = None


def _clone_model(m):
    import copy
    return copy.deepcopy(m)


@pytest.mark.parametrize("variant", sorted(synth.SGD_VARIANTS))
def test_sgd_shim_equals_torch_cpu_sgd(default_ctx, variant):
    """b200ddp.SGD on the GPU vs torch.optim.SGD on the CPU (the reference's
    optimizer, codes/task2/model.py:131) fed identical gradients: bit-exact."""
    hp = synth.SGD_VARIANTS[variant]
    torch.manual_seed(0)
    cpu = ref_ddp.LabNet()
    gpu = _clone_model(cpu).cuda()
    o_ref, o_new = torch.optim.SGD(cpu.parameters(), **hp), b200ddp.SGD(gpu.parameters(), **hp)
    for s in range(4):
        for t, (pc, pg) in enumerate(zip(cpu.parameters(), gpu.parameters())):
            g = torch.from_numpy(synth.dense_grad(pc.numel(), 0, s, t)).view_as(pc)
            pc.grad, pg.grad = g.clone(), g.cuda()
        b200ddp.allreduce_average_gradients(gpu)          # one rank: leaves p.grad as it is
        o_ref.step()
        o_new.step()
    for pc, pg in zip(cpu.parameters(), gpu.parameters()):
        assert torch.equal(pc.detach(), pg.detach().cpu())
        if hp.get("momentum", 0):
            assert torch.equal(o_ref.state[pc]["momentum_buffer"], o_new.state[pg]["momentum_buffer"].cpu())
    assert o_new.launches == 4                            # one kernel per step


def test_sgd_shim_param_groups_state_dict_and_missing_grads(default_ctx):
    torch.manual_seed(1)
    a, b, c = (torch.nn.Parameter(torch.randn(n, device="cuda")) for n in (1000, 33, 5))
    ref = [torch.nn.Parameter(t.detach().cpu().clone()) for t in (a, b, c)]
    groups = lambda ps: [{"params": [ps[0]], "lr": 0.1}, {"params": [ps[1], ps[2]], "lr": 0.01, "momentum": 0.5}]
    o_new, o_ref = b200ddp.SGD(groups([a, b, c]), lr=1.0, momentum=0.9), torch.optim.SGD(groups(ref), lr=1.0, momentum=0.9)
    for s in range(3):
        for t, (pg, pc) in enumerate(zip((a, b, c), ref)):
            if t == 2 and s == 0:
                pg.grad = pc.grad = None             # a parameter without gradient is skipped, as in torch
                continue
            g = torch.from_numpy(synth.dense_grad(pc.numel(), 0, s, t))
            pg.grad, pc.grad = g.cuda(), g.clone()
        o_new.step()
        o_ref.step()
    for pg, pc in zip((a, b, c), ref):
        assert torch.equal(pg.detach().cpu(), pc.detach())
    sd = o_new.state_dict()
    assert sd["param_groups"][1]["momentum"] == 0.5 and len(sd["state"]) == 3
    o2 = b200ddp.SGD(groups([a, b, c]), lr=1.0, momentum=0.9)
    o2.load_state_dict(sd)
    assert torch.equal(o2.state[b]["momentum_buffer"], o_new.state[b]["momentum_buffer"])
    o_new.zero_grad()
    assert a.grad is None


def test_constructor_rejects_what_torch_rejects():
    p = [torch.nn.Parameter(torch.zeros(1))]
    for kw in (dict(lr=-1.0), dict(momentum=-0.1), dict(weight_decay=-1.0), dict(nesterov=True),
               dict(nesterov=True, momentum=0.9, dampening=0.1)):
        with pytest.raises(ValueError):
            b200ddp.SGD(p, **({"lr": 0.1} | kw))
        with pytest.raises(ValueError):
            torch.optim.SGD(p, **({"lr": 0.1} | kw))


def test_dist_utils_names_on_one_rank(default_ctx):
    du = b200ddp.dist_utils
    assert du.get_world_size() == 1 and du.get_local_rank() == 0      # codes/task2/dist_utils.py:18-30
    m = ref_ddp.LabNet().cuda()
    du.init_parameters(m)                                             # no-op on one rank
    with pytest.raises(AttributeError):
        m0 = ref_ddp.LabNet().cuda()
        keep = b200ddp.SGD(m0.parameters(), lr=0.1)
        du.enable_fused_step(m0, keep)
        du.allreduce_average_gradients(m0)                            # no grads yet: same failure mode as the reference
    for p in m.parameters():
        p.grad = torch.ones_like(p)
    du.average_gradients(m)
    du.allgather_average_gradients(m)
    assert all(torch.equal(p.grad, torch.ones_like(p)) for p in m.parameters())


def test_training_loop_matches_torch_sgd_on_gpu(default_ctx):
    """The reference's hot loop on the GPU, our step vs torch's own CUDA SGD: the
    same forward/backward kernels feed both, so parameters agree to rel 1e-6."""
    torch.manual_seed(0)
    m1 = ref_ddp.LabNet().cuda()
    m2 = _clone_model(m1)
    o1, o2 = b200ddp.SGD(m1.parameters(), lr=0.01, momentum=0.9), torch.optim.SGD(m2.parameters(), lr=0.01, momentum=0.9)
    loss_fn = torch.nn.CrossEntropyLoss()
    for s in range(5):
        x, y = ref_ddp.synthetic_batch(0, s, 32, (1, 28, 28), 10)
        x, y = x.cuda(), y.cuda()
        for m, o in ((m1, o1), (m2, o2)):
            loss = loss_fn(m(x), y)
            o.zero_grad()
            loss.backward()
            if m is m1:
                b200ddp.allreduce_average_gradients(m)
            o.step()
    for a, b in zip(m1.parameters(), m2.parameters()):
        assert torch.allclose(a, b, rtol=1e-6, atol=1e-8)


def test_comm_hook_on_one_rank(default_ctx):
    class FakeBucket:                      # GradBucket surface used by the hook (torch/_C/_distributed_c10d.pyi:34-40)
        def __init__(self, t):
            self.t = t

        def buffer(self):
            return self.t

        def index(self):
            return 0

    g = torch.from_numpy(synth.dense_grad(52000, 0, 0, 0)).cuda()
    want = g.clone()
    fut = b200ddp.comm_hook(b200ddp.HookState(default_ctx), FakeBucket(g))
    out = fut.wait()
    assert out.data_ptr() == g.data_ptr() and torch.equal(out, want)


def _mlp():
    torch.manual_seed(3)
    return torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(784, 300), torch.nn.ReLU(),
                               torch.nn.Linear(300, 100), torch.nn.ReLU(), torch.nn.Linear(100, 10)).cuda()


@pytest.mark.parametrize("fused", [False, True])
def test_graphed_train_step_equals_eager_loop(default_ctx, fused):
    """b200ddp.GraphedTrainStep replays the reference's loop body from a CUDA graph;
    it must leave exactly the state the eager loop leaves (linear layers only, so
    forward/backward are run-to-run deterministic and the comparison is bit-exact)."""
    results = []
    for graphed in (False, True):
        model = _mlp()
        opt = b200ddp.SGD(model.parameters(), lr=0.01, momentum=0.9)
        if fused:
            b200ddp.enable_fused_step(model, opt)
        loss_fn = torch.nn.CrossEntropyLoss()
        step = b200ddp.GraphedTrainStep(model, loss_fn, opt, eager_steps=2) if graphed else None
        losses = []
        for s in range(9):
            x, y = ref_ddp.synthetic_batch(0, s, 32, (1, 28, 28), 10)
            if graphed:
                losses.append(float(step(x.pin_memory(), y.pin_memory()).item()))
            else:
                x, y = x.cuda(), y.cuda()
                loss = loss_fn(model(x), y)
                opt.zero_grad()
                loss.backward()
                b200ddp.allreduce_average_gradients(model)
                opt.step()
                losses.append(float(loss.item()))
        if graphed:
            assert step.capture_error is None and step.graph is not None and step.calls == 9
        results.append((losses, [p.detach().clone() for p in model.parameters()],
                        [opt.state[p]["momentum_buffer"].clone() for p in model.parameters()],
                        [p.grad.clone() for p in model.parameters()]))
    (l0, p0, m0, g0), (l1, p1, m1, g1) = results
    assert l0 == l1
    for a, b in zip(p0 + m0 + g0, p1 + m1 + g1):
        assert torch.equal(a, b)


def test_context_options_round_trip_and_reject_bad_values(ctx1):
    """b200ddp_set_option / b200ddp_get_option (include/b200ddp.h, B200DDP_OPT_*)."""
    assert ctx1.get_option("div_mode") == b200ddp._lib.DIV_RECIPROCAL       # the documented default
    assert ctx1.get_option("peer_timeout_ms") == 0                          # wait for ever, like NCCL
    for name, value in (("div_mode", "ieee"), ("pdl", 0), ("nvls_min_bytes", 1 << 20), ("oneshot_max_bytes", 4096),
                        ("peer_timeout_ms", 1500), ("p2p_blocks", 64), ("nvls_depth", 8), ("copy_unroll", 8),
                        ("hbm_blocks_per_sm", 2), ("trace", 1)):
        ctx1.set_option(name, value)
        want = {"ieee": b200ddp._lib.DIV_IEEE}.get(value, value)
        assert ctx1.get_option(name) == want, name
    for name, value in (("div_mode", 7), ("p2p_blocks", 100000), ("nvls_depth", 3), ("copy_unroll", 5),
                        ("hbm_blocks_per_sm", -1)):
        with pytest.raises(b200ddp.B200DDPError):
            ctx1.set_option(name, value)
    ctx1.set_option("trace", 0)
    with pytest.raises(b200ddp.B200DDPError, match="tracing is off"):
        ctx1.trace_read(4)


@pytest.mark.parametrize("unroll", [2, 4, 8])
@pytest.mark.parametrize("blocks_per_sm", [0, 1])
def test_pack_unpack_launch_settings_do_not_change_results(ctx1, unroll, blocks_per_sm):
    """copy_unroll changes the chunking of pack / unpack, hbm_blocks_per_sm their grid: same bytes out."""
    ctx1.set_option("copy_unroll", unroll).set_option("hbm_blocks_per_sm", blocks_per_sm)
    numels = list(synth.RAGGED_NUMELS) + [100_003, 8192 * 3]
    plan = ctx1.plan(numels)
    src = [torch.from_numpy(synth.dense_grad(n, 0, 0, t)).cuda() for t, n in enumerate(numels)]
    out = [torch.full_like(t, float("nan")) for t in src]
    ctx1.pack_scale(plan, [t.data_ptr() for t in src], ctx1.bucket_ptr(0), 0.25)
    ctx1.unpack(plan, ctx1.bucket_ptr(0), [t.data_ptr() for t in out], 1.0)
    for a, b in zip(src, out):
        assert torch.equal(a * 0.25, b)
    # and the update kernel under the grid cap
    p = [torch.ones_like(t) for t in src]
    ctx1.unpack_sgd(plan, ctx1.bucket_ptr(0), [t.data_ptr() for t in p], None, None, b200ddp.sgd_args(lr=1.0))
    for a, b in zip(src, p):
        assert torch.equal(b, 1.0 - a * 0.25)
