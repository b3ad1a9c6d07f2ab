"""Host orchestration of the sync step for N > 1, on the CPU: GradSync driven against a
recording stand-in for the context (no kernels run - this checks WHICH calls are made, with which
scale / divisor / algorithm / offsets, for every world size)."""
import types

import pytest
import torch

from b200ddp import _lib
from b200ddp.engine import GradSync, padded_numel


class FakeTensor:
    """Just enough of a CUDA tensor for the engine's checks."""
    _next = 0x7000_0000_0000

    def __init__(self, numel, with_grad=True):
        self._n = numel
        FakeTensor._next += 1 << 24
        self._ptr = FakeTensor._next
        self.dtype, self.is_cuda = torch.float32, True
        self.grad = FakeTensor(numel, with_grad=False) if with_grad else None

    def numel(self):
        return self._n

    def is_contiguous(self):
        return True

    def data_ptr(self):
        return self._ptr


class FakeCtx:
    def __init__(self, world, staging_bytes=64 << 20, oneshot_max=256 << 10):
        self.world, self.rank, self.staging_bytes = world, 0, staging_bytes
        self.staging_ptr = 0x1000_0000
        self.oneshot_max = oneshot_max
        self.calls = []

    def plan(self, numels, dtype):
        elems = sum(padded_numel(n) for n in numels)
        esize = 2 if dtype == _lib.BF16 else 4
        return types.SimpleNamespace(numels=list(numels), bucket_elems=elems, bucket_bytes=elems * esize, dtype=dtype,
                                     close=lambda: None)

    def bucket_ptr(self, off):
        return self.staging_ptr + off

    def allreduce_pick(self, nelem, dtype):
        nbytes = nelem * (2 if dtype == _lib.BF16 else 4)
        return _lib.ALGO_ONESHOT if nbytes <= self.oneshot_max else _lib.ALGO_TWOSHOT

    def pack_scale(self, plan, ptrs, bucket, scale, stream=None):
        self.calls.append(("pack", len(list(ptrs)), bucket - self.staging_ptr, scale))

    def allreduce(self, off, nelem, dtype, algo, stream=None):
        self.calls.append(("allreduce", off, nelem, algo))
        return off + (self.staging_bytes if algo == _lib.ALGO_ONESHOT else 0)     # one-shot result is out of place

    def unpack(self, plan, bucket, ptrs, divisor, stream=None):
        self.calls.append(("unpack", bucket - self.staging_ptr, divisor))

    def unpack_sgd(self, plan, bucket, ps, ms, gs, args, stream=None):
        self.calls.append(("unpack_sgd", bucket - self.staging_ptr, args.grad_divisor, args.first_step, gs is not None))

    def allreduce_sgd(self, plan, off, ps, ms, gs, args):
        self.calls.append(("allreduce_sgd", off, args.grad_divisor, args.first_step, gs is not None))


def _params(numels):
    return [FakeTensor(n) for n in numels]


@pytest.mark.parametrize("world", [2, 3, 4, 5, 8])
def test_strict_sync_scale_policy_and_call_order(world):
    """Power-of-two worlds pre-scale by 1/N in the pack and do not divide; others pack unscaled
    and divide by N after the sum (the reference's order, codes/task2/dist_utils.py:41-42)."""
    ctx = FakeCtx(world)
    eng = GradSync(_params([300_000, 17]), ctx, algo="auto")
    eng.sync()
    kinds = [c[0] for c in ctx.calls]
    assert kinds == ["pack", "allreduce", "unpack"] and eng.launches == 3
    pow2 = world & (world - 1) == 0
    assert ctx.calls[0][3] == (1.0 / world if pow2 else 1.0)
    assert ctx.calls[2][2] == (1.0 if pow2 else float(world))
    assert ctx.calls[1][3] == _lib.ALGO_TWOSHOT and ctx.calls[2][1] == 0            # in place


def test_strict_sync_reads_the_one_shot_result_from_the_second_half():
    ctx = FakeCtx(4)
    eng = GradSync(_params([1000, 10]), ctx, algo="auto")
    eng.sync()
    assert ctx.calls[1][3] == _lib.ALGO_ONESHOT
    assert ctx.calls[2][1] == ctx.staging_bytes                                       # unpack from the result half


def test_one_rank_is_a_no_op():
    ctx = FakeCtx(1)
    eng = GradSync(_params([1000]), ctx)
    eng.sync()
    assert ctx.calls == [] and eng.launches == 0


@pytest.mark.parametrize("numel,expect", [(1000, ["pack", "allreduce_sgd"]), (300_000, ["pack", "allreduce", "unpack_sgd"])])
def test_fused_step_uses_the_fused_kernel_only_for_small_buckets(numel, expect):
    ctx = FakeCtx(8)
    eng = GradSync(_params([numel]), ctx, algo="auto")
    moms = [FakeTensor(numel, with_grad=False)]
    eng.sync_and_step(moms, dict(lr=0.01, momentum=0.9), first_step=True)
    assert [c[0] for c in ctx.calls] == expect and eng.launches == len(expect)
    last = ctx.calls[-1]
    assert last[2] == 1.0 and last[3] == 1 and last[4] is True          # 8 ranks: no divisor; first step; p.grad written


def test_fused_step_for_an_odd_world_divides_in_the_update():
    ctx = FakeCtx(3)
    eng = GradSync(_params([300_000]), ctx, algo="twoshot")
    eng.sync_and_step([FakeTensor(300_000, with_grad=False)], dict(lr=0.01, momentum=0.9), first_step=False, write_grads=False)
    assert ctx.calls[0][3] == 1.0 and ctx.calls[-1][0] == "unpack_sgd"
    assert ctx.calls[-1][2] == 3.0 and ctx.calls[-1][3] == 0 and ctx.calls[-1][4] is False


def test_bf16_buckets_always_prescale():
    ctx = FakeCtx(3)
    eng = GradSync(_params([300_000]), ctx, bucket_dtype="bf16", algo="twoshot")
    eng.sync()
    assert ctx.calls[0][3] == pytest.approx(1.0 / 3) and ctx.calls[-1][2] == 1.0


def test_model_larger_than_staging_is_cut_into_buckets_with_reused_offsets():
    ctx = FakeCtx(2, staging_bytes=1 << 20)
    eng = GradSync(_params([150_000] * 5), ctx, algo="twoshot")           # 600 KB each: one per bucket
    eng.sync()
    packs = [c for c in ctx.calls if c[0] == "pack"]
    assert len(packs) == 5 and {c[2] for c in packs} == {0}                # every bucket restarts at offset 0
    assert [c[0] for c in ctx.calls] == ["pack", "allreduce", "unpack"] * 5


def test_missing_gradient_raises_like_the_reference():
    ctx = FakeCtx(2)
    ps = _params([100, 100])
    ps[1].grad = None
    with pytest.raises(AttributeError):
        GradSync(ps, ctx).sync()


def test_reverse_split_puts_the_tail_of_the_list_first():
    """Overlap mode cuts buckets from the LAST parameter backwards (the order backward produces
    gradients in): bucket 0 is the tail, the last bucket holds parameter 0, every tensor is in
    exactly one bucket, and each bucket keeps list order inside."""
    from b200ddp.engine import split_buckets_reverse
    numels = [9408, 64, 64, 36864, 64, 64, 2_359_296, 512, 512, 5120, 10]
    layout = split_buckets_reverse(numels, 4, 4 << 20, 64 << 20)
    assert layout[0][1] == len(numels) and layout[-1][0] == 0
    covered = []
    for lo, hi, off, elems in layout:
        assert lo < hi and elems == sum(padded_numel(n) for n in numels[lo:hi]) and off % 256 == 0
        covered = list(range(lo, hi)) + covered
    assert covered == list(range(len(numels)))
    # offsets do not overlap
    spans = sorted((off, off + elems * 4) for _, _, off, elems in layout)
    assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:]))
