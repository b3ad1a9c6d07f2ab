"""K3 with the lab's own update rules (codes/task1/pytorch/MyOptimizer.py:18-43) against
  * the reference classes' CPU results (tests/golden/opt_*.npz, oracle/make_golden.py),
  * the reference's three statements run by torch ON THIS GPU (oracle.ref_ddp.adam_step_torch), and
  * the numpy restatement with IEEE sqrt / division.
GD is bit-exact against all three.  Adam is bit-exact against torch-on-CUDA and numpy; against the
CPU golden its parameters are held to rel <= 1e-6, because ATen's CPU sqrt is not correctly rounded
(oracle/ref_ddp.py); m and v involve no sqrt and are bit-exact there too."""
import os

import numpy as np
import pytest
import torch

import b200ddp
from helpers import dev, host, ptrs, to_dev
from oracle import ref_ddp, synth
from oracle.make_golden import lab_optimizer_grad

pytestmark = pytest.mark.gpu
NUMELS = synth.RAGGED_NUMELS


def _kw(name, hp):
    return dict(lr=hp[0]) if name == "gd" else dict(lr=hp[0], beta1=hp[1], beta2=hp[2], eps=hp[3])


@pytest.mark.parametrize("mode", ["direct", "bucket", "bucket_nograd"])
@pytest.mark.parametrize("name", ["gd", "adam", "adam_b"])
def test_lab_rules_through_the_c_abi(ctx1, golden_dir, name, mode):
    g = dict(np.load(os.path.join(golden_dir, f"opt_{name}.npz")))
    steps, kw = int(g["steps"]), _kw(name, list(g["hp"]))
    rule = "gd" if name == "gd" else "adam"
    plan = ctx1.plan(NUMELS)
    p = to_dev([synth.init_param(n, t) for t, n in enumerate(NUMELS)])
    m = [torch.zeros_like(t) for t in p] if rule == "adam" else None
    v = [torch.zeros_like(t) for t in p] if rule == "adam" else None
    gout = [torch.full_like(t, float("nan")) for t in p]
    # the reference's statements, run by torch on this GPU
    pt = [t.clone() for t in p]
    mt, vt = [torch.zeros_like(t) for t in p], [torch.zeros_like(t) for t in p]
    for s in range(steps):
        grads = to_dev([lab_optimizer_grad(n, s, t) for t, n in enumerate(NUMELS)])
        args = b200ddp.opt_args(rule, **kw)
        if mode == "direct":
            ctx1.opt_step(plan, ptrs(grads), ptrs(p), ptrs(m) if m else None, ptrs(v) if v else None, args)
        else:
            ctx1.pack_scale(plan, ptrs(grads), ctx1.bucket_ptr(0), 1.0)
            ctx1.unpack_opt(plan, ctx1.bucket_ptr(0), ptrs(p), ptrs(m) if m else None, ptrs(v) if v else None,
                            ptrs(gout) if mode == "bucket" else None, args)
            if mode == "bucket":
                assert all(torch.equal(a, b) for a, b in zip(gout, grads))
        for t in range(len(NUMELS)):
            if rule == "gd":
                pt[t] = ref_ddp.gd_step_torch(pt[t], grads[t], **kw)
            else:
                pt[t], mt[t], vt[t] = ref_ddp.adam_step_torch(pt[t], grads[t], mt[t], vt[t], **kw)
    got_p = np.concatenate(host(p))
    assert np.array_equal(got_p, np.concatenate(host(pt))), "differs from the reference's statements run by torch on CUDA"
    if rule == "gd":
        assert np.array_equal(got_p, g["p"])
    else:
        assert np.array_equal(np.concatenate(host(m)), g["m"]) and np.array_equal(np.concatenate(host(v)), g["v"])
        assert np.array_equal(np.concatenate(host(m)), np.concatenate(host(mt)))
        assert np.array_equal(np.concatenate(host(v)), np.concatenate(host(vt)))
        assert np.max(np.abs(got_p - g["p"])) <= 1e-6 * np.max(np.abs(g["p"]))


def test_adam_matches_numpy_ieee_restatement(ctx1):
    n = 70_001
    rng = np.random.default_rng(5)
    p0 = rng.standard_normal(n).astype(np.float32)
    plan = ctx1.plan((n,))
    p, m, v = to_dev([p0])[0], torch.zeros(n, device=dev()), torch.zeros(n, device=dev())
    pn, mn, vn = p0.copy(), np.zeros(n, np.float32), np.zeros(n, np.float32)
    kw = dict(lr=0.001, beta1=0.9, beta2=0.999, eps=1e-8)
    for s in range(3):
        g = (rng.standard_normal(n) * 10.0 ** rng.integers(-8, 2, n)).astype(np.float32)
        g[::7] = 0.0
        gd = to_dev([g])[0]
        ctx1.opt_step(plan, [gd.data_ptr()], [p.data_ptr()], [m.data_ptr()], [v.data_ptr()], b200ddp.opt_args("adam", **kw))
        pn, mn, vn = ref_ddp.adam_step_numpy(pn, g, mn, vn, **kw)
    for got, want in ((p, pn), (m, mn), (v, vn)):
        assert np.array_equal(got.cpu().numpy().view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("cls", ["gd", "adam"])
def test_optimizer_classes_keep_the_labs_interface(built_lib, golden_dir, cls):
    """Constructor, attributes, step() and zero_grad() of codes/task1/pytorch/MyOptimizer.py."""
    torch.cuda.set_device(0)
    b200ddp.reset_default_context()
    g = dict(np.load(os.path.join(golden_dir, f"opt_{cls}.npz")))
    hp = list(g["hp"])
    params = [torch.nn.Parameter(t) for t in to_dev([synth.init_param(n, t) for t, n in enumerate(NUMELS)])]
    if cls == "gd":
        opt = b200ddp.GdOptimizer(params, lr=hp[0])
    else:
        opt = b200ddp.AdamOptimizer(params, lr=hp[0], b1=hp[1], b2=hp[2], epsilon=hp[3])
        assert (opt.beta1, opt.beta2, opt.epsilon) == (hp[1], hp[2], hp[3]) and len(opt.momentums) == len(params)
    assert opt.lr == hp[0] and [id(a) for a in opt.params] == [id(a) for a in params]
    for s in range(int(g["steps"])):
        for t, q in enumerate(params):
            q.grad = to_dev([lab_optimizer_grad(q.numel(), s, t)])[0]
        opt.step()
    got = np.concatenate([q.detach().cpu().numpy() for q in params])
    if cls == "gd":
        assert np.array_equal(got, g["p"])
    else:
        assert np.max(np.abs(got - g["p"])) <= 1e-6 * np.max(np.abs(g["p"]))
        assert np.array_equal(np.concatenate([t.cpu().numpy() for t in opt.momentums]), g["m"])
    opt.zero_grad()
    assert all(float(q.grad.abs().max()) == 0.0 for q in params if q.grad.numel())
    params[3].grad = None                      # a parameter without gradient is skipped (MyOptimizer.py:22,37)
    before = params[3].detach().clone()
    opt.step()
    assert torch.equal(params[3].detach(), before)
    b200ddp.reset_default_context()
