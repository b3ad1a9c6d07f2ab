"""Parity of the CUDA kernels (through the C ABI) with the CPU oracle.
Bit-exact unless a tolerance is stated in the test."""
import os

import numpy as np
import pytest
import torch

import b200ddp
from b200ddp import _lib
from oracle import ref_ddp, synth

from helpers import EmuWorld, dev, from_bucket, host, ptrs, read_bucket, to_dev

pytestmark = pytest.mark.gpu

RAGGED = (0,) + synth.RAGGED_NUMELS + (2048, 2049, 6143, 6144, 70001)


def _grads(numels, rank=0, step=0, kind=synth.dense_grad):
    return [kind(n, rank, step, t) for t, n in enumerate(numels)]


# ----------------------------------------------------------------------------- K1 pack
@pytest.mark.parametrize("scale", [1.0, 0.5, 1.0 / 3.0, 0.125])
def test_pack_fp32_matches_oracle(ctx1, scale):
    plan = ctx1.plan(RAGGED)
    g = _grads(RAGGED)
    gd = to_dev(g)
    ctx1.pack_scale(plan, ptrs(gd), ctx1.bucket_ptr(0), scale)
    flat = read_bucket(ctx1, 0, plan.bucket_elems).numpy()
    for got, ref in zip(from_bucket(flat, plan), g):
        assert np.array_equal(got, ref * np.float32(scale))
    # padding between tensors is never written and stays zero
    mask = np.ones(plan.bucket_elems, bool)
    for o, n in zip(plan.offsets, plan.numels):
        mask[o:o + n] = False
    assert not flat[mask].any()
    assert all(o % 32 == 0 for o in plan.offsets)


def test_pack_bf16_matches_torch_cast(ctx1):
    plan = ctx1.plan(RAGGED, "bf16")
    g = _grads(RAGGED)
    gd = to_dev(g)
    ctx1.pack_scale(plan, ptrs(gd), ctx1.bucket_ptr(0), 0.25)
    flat = read_bucket(ctx1, 0, plan.bucket_elems, torch.bfloat16)
    for o, n, ref in zip(plan.offsets, plan.numels, g):
        want = (torch.from_numpy(ref) * 0.25).to(torch.bfloat16)      # round-to-nearest-even, as the kernel
        assert torch.equal(flat[o:o + n], want)


def test_pack_misaligned_tensors_take_the_scalar_path(ctx1):
    numels = (5000, 2048, 4099)
    plan = ctx1.plan(numels)
    g = _grads(numels)
    backing = [torch.zeros(n + 3, device=dev()) for n in numels]
    views = []
    for b, a, shift in zip(backing, g, (1, 2, 3)):          # 4-, 8-, 12-byte offsets: not 16-byte aligned
        v = b[shift:shift + len(a)]
        v.copy_(torch.from_numpy(a))
        assert v.data_ptr() % 16 != 0
        views.append(v)
    ctx1.pack_scale(plan, ptrs(views), ctx1.bucket_ptr(0), 0.5)
    flat = read_bucket(ctx1, 0, plan.bucket_elems).numpy()
    for got, ref in zip(from_bucket(flat, plan), g):
        assert np.array_equal(got, ref * np.float32(0.5))


# ----------------------------------------------------------------------------- unpack
RULES = {"ieee": "cpu", "reciprocal": "cuda"}      # library option -> oracle rule


@pytest.mark.parametrize("rule", sorted(RULES))
@pytest.mark.parametrize("divisor", [1.0, 2.0, 3.0, 7.0])
def test_pack_unpack_divides_like_the_reference(ctx1, divisor, rule):
    ctx1.set_option("div_mode", rule)
    plan = ctx1.plan(RAGGED)
    g = _grads(RAGGED)
    gd = to_dev(g)
    out = [torch.full_like(t, float("nan")) for t in gd]
    ctx1.pack_scale(plan, ptrs(gd), ctx1.bucket_ptr(0), 1.0)
    ctx1.unpack(plan, ctx1.bucket_ptr(0), ptrs(out), divisor)
    for got, ref in zip(host(out), g):
        assert np.array_equal(got, ref_ddp.divide_by_world(ref, divisor, RULES[rule]))   # `param.grad.data /= world_size`


@pytest.mark.parametrize("n", [3, 5, 6, 7, 2, 8])
def test_reciprocal_division_is_what_torch_cuda_computes(ctx1, n):
    """Pins the "cuda" rule of the oracle and the library's default on the real thing: torch on this
    GPU dividing a tensor in place by a Python int, as codes/task2/dist_utils.py:42 does."""
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(200_003) * 10.0 ** rng.integers(-30, 30, 200_003)).astype(np.float32)
    t = torch.from_numpy(x).to(dev())
    ref = t.clone()
    ref /= n
    ref = ref.cpu().numpy()
    assert np.array_equal(ref.view(np.uint32), ref_ddp.divide_by_world(x, n, "cuda").view(np.uint32))
    assert ctx1.get_option("div_mode") == b200ddp._lib.DIV_RECIPROCAL          # the default
    plan = ctx1.plan((x.size,))
    out = torch.empty_like(t)
    ctx1.pack_scale(plan, [t.data_ptr()], ctx1.bucket_ptr(0), 1.0)
    ctx1.unpack(plan, ctx1.bucket_ptr(0), [out.data_ptr()], float(n))
    assert np.array_equal(out.cpu().numpy().view(np.uint32), ref.view(np.uint32))


# ----------------------------------------------------------------------------- K3 update
def _torch_sgd_reference(numels, hp, steps, grad_fn):
    params = [torch.nn.Parameter(torch.from_numpy(synth.init_param(n, t))) for t, n in enumerate(numels)]
    opt = torch.optim.SGD(params, **hp)
    for s in range(steps):
        for t, p in enumerate(params):
            p.grad = torch.from_numpy(grad_fn(p.numel(), s, t))
        opt.step()
    p = [q.detach().numpy() for q in params]
    m = [opt.state[q]["momentum_buffer"].numpy() for q in params] if hp.get("momentum", 0) else None
    return p, m


@pytest.mark.parametrize("variant", sorted(synth.SGD_VARIANTS))
@pytest.mark.parametrize("mode", ["bucket", "bucket_nograd", "direct"])
def test_update_matches_torch_sgd(ctx1, variant, mode):
    hp = synth.SGD_VARIANTS[variant]
    numels = RAGGED[1:]
    steps = 3
    grad_fn = lambda n, s, t: synth.dense_grad(n, 0, s, t)
    p_ref, m_ref = _torch_sgd_reference(numels, hp, steps, grad_fn)
    plan = ctx1.plan(numels)
    p = to_dev([synth.init_param(n, t) for t, n in enumerate(numels)])
    m = [torch.full_like(t, float("nan")) for t in p] if hp.get("momentum", 0) else None
    gout = [torch.full_like(t, float("nan")) for t in p]
    for s in range(steps):
        g = to_dev([grad_fn(n, s, t) for t, n in enumerate(numels)])
        args = b200ddp.sgd_args(first_step=(s == 0), **hp)
        if mode == "direct":
            ctx1.sgd_step(plan, ptrs(g), ptrs(p), ptrs(m) if m else None, args)
        else:
            ctx1.pack_scale(plan, ptrs(g), ctx1.bucket_ptr(0), 1.0)
            ctx1.unpack_sgd(plan, ctx1.bucket_ptr(0), ptrs(p), ptrs(m) if m else None,
                            ptrs(gout) if mode == "bucket" else None, args)
            if mode == "bucket":
                for got, ref in zip(host(gout), host(g)):
                    assert np.array_equal(got, ref)
    for got, ref in zip(host(p), p_ref):
        assert np.array_equal(got, ref)
    if m:
        for got, ref in zip(host(m), m_ref):
            assert np.array_equal(got, ref)


@pytest.mark.parametrize("rule", sorted(RULES))
def test_update_with_divisor_matches_sum_then_divide(ctx1, rule):
    """Non-power-of-two worlds: the bucket holds the SUM and K3 divides, as
    codes/task2/dist_utils.py:42 does."""
    ctx1.set_option("div_mode", rule)
    numels, hp = (1000, 4099), dict(lr=0.01, momentum=0.9)
    plan = ctx1.plan(numels)
    gsum = _grads(numels)
    p0 = [synth.init_param(n, t) for t, n in enumerate(numels)]
    p, m, gout = to_dev(p0), [torch.empty(n, device=dev()) for n in numels], [torch.empty(n, device=dev()) for n in numels]
    ctx1.pack_scale(plan, ptrs(to_dev(gsum)), ctx1.bucket_ptr(0), 1.0)
    ctx1.unpack_sgd(plan, ctx1.bucket_ptr(0), ptrs(p), ptrs(m), ptrs(gout),
                    b200ddp.sgd_args(first_step=True, grad_divisor=3.0, **hp))
    for t in range(len(numels)):
        avg = ref_ddp.divide_by_world(gsum[t], 3, RULES[rule])
        pr, mr = ref_ddp.sgd_step_numpy(p0[t], avg, None, **hp)
        assert np.array_equal(host(gout)[t], avg) and np.array_equal(host(p)[t], pr) and np.array_equal(host(m)[t], mr)


def test_update_from_bf16_bucket(ctx1):
    numels, hp = RAGGED[1:], dict(lr=0.01, momentum=0.9)
    plan = ctx1.plan(numels, "bf16")
    g = _grads(numels)
    p0 = [synth.init_param(n, t) for t, n in enumerate(numels)]
    p, m = to_dev(p0), [torch.empty(n, device=dev()) for n in numels]
    gout = [torch.empty(n, device=dev()) for n in numels]
    ctx1.pack_scale(plan, ptrs(to_dev(g)), ctx1.bucket_ptr(0), 1.0)
    ctx1.unpack_sgd(plan, ctx1.bucket_ptr(0), ptrs(p), ptrs(m), ptrs(gout), b200ddp.sgd_args(first_step=True, **hp))
    for t in range(len(numels)):
        gb = torch.from_numpy(g[t]).to(torch.bfloat16).to(torch.float32).numpy()    # what the wire carried
        pr, mr = ref_ddp.sgd_step_numpy(p0[t], gb, None, **hp)
        assert np.array_equal(host(gout)[t], gb) and np.array_equal(host(p)[t], pr) and np.array_equal(host(m)[t], mr)


# ----------------------------------------------------------------------------- golden: N ranks emulated on one GPU
def _run_emulated(ws, numels, local_grads, p0, hp, steps, algo, dtype="fp32", fused=False, rule=None):
    """local_grads[s][r] -> list of per-tensor arrays.  Returns per-rank (p, m, avg) after `steps`."""
    w = EmuWorld(ws)
    try:
        if rule is not None:
            for c in w.ctxs:
                c.set_option("div_mode", rule)
        plans = [c.plan(numels, dtype) for c in w.ctxs]
        P = [to_dev(p0) for _ in range(ws)]
        M = [[torch.full((n,), float("nan"), device=dev()) for n in numels] for _ in range(ws)]
        G = [[torch.full((n,), float("nan"), device=dev()) for n in numels] for _ in range(ws)]
        pow2 = ws & (ws - 1) == 0
        pre, div = (1.0 / ws, 1.0) if (pow2 or dtype == "bf16") else (1.0, float(ws))
        has_m = hp.get("momentum", 0) != 0
        for s in range(steps):
            args = b200ddp.sgd_args(first_step=(s == 0), grad_divisor=div, **hp)
            for r, c in enumerate(w.ctxs):
                c.pack_scale(plans[r], ptrs(to_dev(local_grads[s][r])), c.bucket_ptr(0), pre)
            if fused:
                w.allreduce_sgd(plans, 0, P, M if has_m else None, G, args)
            else:
                res = w.allreduce(0, plans[0].bucket_elems, dtype, algo)
                for r, c in enumerate(w.ctxs):
                    c.unpack_sgd(plans[r], c.bucket_ptr(res), ptrs(P[r]), ptrs(M[r]) if has_m else None, ptrs(G[r]), args)
        torch.cuda.synchronize()
        return [(np.concatenate(host(P[r])), np.concatenate(host(M[r])) if has_m else None, np.concatenate(host(G[r])))
                for r in range(ws)]
    finally:
        w.close()


GRID_CASES = [("A", 2), ("A", 4), ("A", 8), ("B", 2), ("C", 2), ("D", 2)]


@pytest.mark.parametrize("variant,ws", GRID_CASES)
@pytest.mark.parametrize("algo", ["oneshot", "twoshot", "fused"])
def test_golden_grid_bit_exact_at_any_world(golden_dir, variant, ws, algo):
    """Injected grid gradients: sums are exact in any order, so the whole step must
    match the reference's gloo run bit for bit at 2, 4 and 8 ranks."""
    g = dict(np.load(os.path.join(golden_dir, f"ddp_grid_{variant}_ws{ws}.npz")))
    numels, hp, steps = synth.RAGGED_NUMELS, synth.SGD_VARIANTS[variant], int(g["steps"])
    local = [[[synth.grid_grad(n, r, s, t) for t, n in enumerate(numels)] for r in range(ws)] for s in range(steps)]
    p0 = [synth.init_param(n, t) for t, n in enumerate(numels)]
    out = _run_emulated(ws, numels, local, p0, hp, steps, algo if algo != "fused" else "oneshot", fused=(algo == "fused"))
    for r in range(ws):
        p, m, avg = out[r]
        assert np.array_equal(avg, g["avg"]), f"rank {r} averaged gradient"
        assert np.array_equal(p, g["p"]), f"rank {r} parameters"
        if m is not None:
            assert np.array_equal(m, g["m"]), f"rank {r} momentum"


@pytest.mark.parametrize("ws", [2, 4])
@pytest.mark.parametrize("algo", ["oneshot", "twoshot", "fused"])
def test_golden_labnet_reference_run(golden_dir, ws, algo):
    """The reference's own hot loop on the lab CNN (codes/task2/model.py:49-63):
    its recorded local gradients in, its averaged gradients / parameters /
    momentum out.  Bit-exact at 2 ranks; at 4 ranks within the stated fp32 bound
    rel <= 1e-5 (inf-norm), the summation order being gloo's vs rank order."""
    g = dict(np.load(os.path.join(golden_dir, f"ddp_labnet_ws{ws}.npz")))
    numels, steps = ref_ddp.LABNET_NUMELS, int(g["steps"])
    splits = np.cumsum(numels)[:-1]
    local = [[np.split(g[f"local_r{r}_s{s}"], splits) for r in range(ws)] for s in range(steps)]
    out = _run_emulated(ws, numels, local, np.split(g["p0"], splits), dict(lr=0.01, momentum=0.9), steps,
                        algo if algo != "fused" else "oneshot", fused=(algo == "fused"))
    for r in range(ws):
        p, m, avg = out[r]
        if ws == 2:
            # NOTE: momentum feeds back, so parameters are compared after the reference's own trajectory
            # diverges only by the summation order - none at 2 ranks.
            assert np.array_equal(avg, g[f"avg_s{steps - 1}"]) and np.array_equal(p, g["p_final"]) and np.array_equal(m, g["m_final"])
        else:
            for got, ref in ((avg, g[f"avg_s{steps - 1}"]), (p, g["p_final"]), (m, g["m_final"])):
                assert np.max(np.abs(got - ref)) <= 1e-5 * np.max(np.abs(ref))
        assert np.array_equal(p, out[0][0]), "ranks must hold identical parameters"


@pytest.mark.parametrize("ws", [4, 8])
@pytest.mark.parametrize("algo", ["oneshot", "twoshot"])
def test_golden_dense_tolerance(golden_dir, ws, algo):
    g = dict(np.load(os.path.join(golden_dir, f"ddp_dense_A_ws{ws}.npz")))
    numels, hp, steps = synth.RAGGED_NUMELS, synth.SGD_VARIANTS["A"], int(g["steps"])
    local = [[[synth.dense_grad(n, r, s, t) for t, n in enumerate(numels)] for r in range(ws)] for s in range(steps)]
    p0 = [synth.init_param(n, t) for t, n in enumerate(numels)]
    out = _run_emulated(ws, numels, local, p0, hp, steps, algo)
    for r in range(ws):
        for got, ref in zip(out[r], (g["p"], g["m"], g["avg"])):
            assert np.max(np.abs(got - ref)) <= 1e-5 * np.max(np.abs(ref))      # fp32 bound of the north star
        assert np.array_equal(out[r][0], out[0][0])


@pytest.mark.parametrize("ws", [2, 3, 8])
@pytest.mark.parametrize("algo", ["oneshot", "twoshot"])
def test_bf16_wire_within_stated_bounds(ws, algo):
    """bf16 buckets: fp32 accumulate, one rounding.  Bounds:
    rel <= 2**-7 vs the fp32 oracle, rel <= 2e-2 vs torch's bf16_compress_hook arithmetic."""
    numels = (4099, 1000, 33)
    local = [[[synth.dense_grad(n, r, 0, t) for t, n in enumerate(numels)] for r in range(ws)]]
    p0 = [synth.init_param(n, t) for t, n in enumerate(numels)]
    out = _run_emulated(ws, numels, local, p0, dict(lr=0.01, momentum=0.9), 1, algo, dtype="bf16")
    per_rank = [torch.from_numpy(np.concatenate(local[0][r])) for r in range(ws)]
    fp32 = ref_ddp.average_over_ranks(per_rank).numpy()
    hook = ref_ddp.bf16_wire_average(per_rank).numpy()
    for r in range(ws):
        avg = out[r][2]
        assert np.max(np.abs(avg - fp32)) <= 2.0 ** -7 * np.max(np.abs(fp32))
        assert np.max(np.abs(avg - hook)) <= 2e-2 * np.max(np.abs(hook))
        assert np.array_equal(avg, out[0][2])


@pytest.mark.parametrize("rule", sorted(RULES))
@pytest.mark.parametrize("algo", ["oneshot", "twoshot", "fused"])
@pytest.mark.parametrize("ws", [3, 5, 6, 7])
def test_odd_world_divides_after_the_sum(ws, algo, rule):
    """Non-power-of-two worlds: sum in rank order, THEN divide - the reference's order - with the
    division rounded as the chosen run of the reference rounds it; bit for bit on every rank."""
    numels = (1000, 4099)
    local = [[[synth.dense_grad(n, r, 0, t) for t, n in enumerate(numels)] for r in range(ws)]]
    p0 = [synth.init_param(n, t) for t, n in enumerate(numels)]
    out = _run_emulated(ws, numels, local, p0, dict(lr=0.01), 1, "oneshot" if algo == "fused" else algo,
                        fused=(algo == "fused"), rule=rule)
    want = ref_ddp.average_over_ranks([torch.from_numpy(np.concatenate(local[0][r])) for r in range(ws)],
                                      RULES[rule]).numpy()
    for r in range(ws):
        assert np.array_equal(out[r][2], want)


# ----------------------------------------------------------------------------- full-size properties
def test_full_size_roundtrip_and_linearity(built_lib):
    """ResNet-50-sized parameter list (25.6 M elements): pack -> unpack is the
    identity, and all-reducing g with -g over two emulated ranks gives exactly 0."""
    import torchvision
    numels = [p.numel() for p in torchvision.models.resnet50().parameters()]
    assert sum(numels) == 25_557_032 and len(numels) == 161
    w = EmuWorld(2, staging_bytes=128 << 20)
    try:
        plans = [c.plan(numels) for c in w.ctxs]
        gen = torch.Generator(device=dev()).manual_seed(1)
        g = [torch.randn(n, device=dev(), generator=gen) * 1e-2 for n in numels]
        neg = [-t for t in g]
        back = [torch.empty_like(t) for t in g]
        w.ctxs[0].pack_scale(plans[0], ptrs(g), w.ctxs[0].bucket_ptr(0), 1.0)
        w.ctxs[0].unpack(plans[0], w.ctxs[0].bucket_ptr(0), ptrs(back), 1.0)
        assert all(torch.equal(a, b) for a, b in zip(g, back))
        w.ctxs[1].pack_scale(plans[1], ptrs(neg), w.ctxs[1].bucket_ptr(0), 1.0)
        for algo in ("twoshot", "oneshot"):
            w.ctxs[0].pack_scale(plans[0], ptrs(g), w.ctxs[0].bucket_ptr(0), 1.0)
            w.ctxs[1].pack_scale(plans[1], ptrs(neg), w.ctxs[1].bucket_ptr(0), 1.0)
            res = w.allreduce(0, plans[0].bucket_elems, "fp32", algo)
            for r in range(2):
                assert not read_bucket(w.ctxs[r], res, plans[0].bucket_elems).any()
        # checksum of checksums: sum over ranks of per-tensor sums equals per-tensor sums of the reduced bucket
        g2 = [torch.randn(n, device=dev(), generator=gen) * 1e-2 for n in numels]
        w.ctxs[0].pack_scale(plans[0], ptrs(g), w.ctxs[0].bucket_ptr(0), 1.0)
        w.ctxs[1].pack_scale(plans[1], ptrs(g2), w.ctxs[1].bucket_ptr(0), 1.0)
        res = w.allreduce(0, plans[0].bucket_elems, "fp32", "twoshot")
        flat = read_bucket(w.ctxs[1], res, plans[0].bucket_elems)
        for o, n, a, b in zip(plans[0].offsets, numels, g, g2):
            assert torch.equal(flat[o:o + n], (a + b).cpu())
    finally:
        w.close()


# ----------------------------------------------------------------------------- error behaviour
def test_errors_are_loud(ctx1):
    plan = ctx1.plan((16, 16))
    t = torch.zeros(16, device=dev())
    with pytest.raises(b200ddp.B200DDPError, match="NULL"):
        ctx1.pack_scale(plan, [t.data_ptr(), 0], ctx1.bucket_ptr(0), 1.0)
    with pytest.raises(b200ddp.B200DDPError, match="exceeds"):
        ctx1.allreduce(0, ctx1.staging_bytes, "fp32")
    with pytest.raises(b200ddp.B200DDPError, match="Nesterov"):
        ctx1.sgd_step(plan, ptrs([t, t]), ptrs([t, t]), ptrs([t, t]), b200ddp.sgd_args(0.1, momentum=0.0, nesterov=True))
    other = b200ddp.Context(0, 1, 0, staging_bytes=1 << 20)
    with pytest.raises(b200ddp.B200DDPError, match="another context"):
        other.pack_scale(plan, ptrs([t, t]), other.bucket_ptr(0), 1.0)
    other.close()
    two = b200ddp.Context(0, 2, 0, staging_bytes=1 << 20)
    with pytest.raises(b200ddp.B200DDPError, match="no collective|peer memory"):
        two.allreduce(0, 64, "fp32")
    two.close()


def test_empty_plan_is_a_noop(ctx1):
    plan = ctx1.plan(())
    ctx1.pack_scale(plan, [], ctx1.bucket_ptr(0), 1.0)
    ctx1.unpack_sgd(plan, ctx1.bucket_ptr(0), [], None, None, b200ddp.sgd_args(0.1))
    assert plan.bucket_elems == 0


@pytest.mark.parametrize("divisor", [3.0, 5.0, 6.0, 7.0, 2.5])
def test_division_edge_values_match_ieee(ctx1, divisor):
    """The short division sequence (small integer divisors) and its __fdiv_rn fallback (everything
    else: zeros, denormals, huge values, non-integer divisors) both give the IEEE quotient."""
    ctx1.set_option("div_mode", "ieee")
    rng = np.random.default_rng(3)
    n = 70001
    x = (rng.standard_normal(n) * 10.0 ** rng.integers(-44, 38, n)).astype(np.float32)
    x[:12] = np.float32([0.0, -0.0, 1e-45, -1e-45, 1.1754944e-38, 3.4e38, -3.4e38, 1e-30, 1e30, 1.0, 3.0, 7.0])
    plan = ctx1.plan((n,))
    src, out = to_dev([x])[0], torch.empty(n, device=dev())
    ctx1.pack_scale(plan, [src.data_ptr()], ctx1.bucket_ptr(0), 1.0)
    ctx1.unpack(plan, ctx1.bucket_ptr(0), [out.data_ptr()], divisor)
    got = out.cpu().numpy()
    with np.errstate(all="ignore"):
        want = (x / np.float32(divisor)).astype(np.float32)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))          # bit patterns, so that -0 and denormals count
