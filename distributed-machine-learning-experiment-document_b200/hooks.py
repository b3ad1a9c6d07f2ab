"""DDP communication hook: the same kernels behind
`DistributedDataParallel.register_comm_hook` (torch/nn/parallel/distributed.py:1987-2067).

    ddp = DistributedDataParallel(model, device_ids=[dev])
    ddp.register_comm_hook(b200ddp.HookState(), b200ddp.comm_hook)

The reference itself never wraps its model in DDP (it averages by hand,
codes/task2/dist_utils.py:39-42); this entry point exists because the north
star names it.  Contract kept: the future resolves to a tensor shaped like
bucket.buffer(); gradients arrive un-divided and leave averaged.
"""
import torch
import torch.distributed as dist

from . import _lib
from .runtime import Context, default_context, dtype_code


class HookState:
    def __init__(self, ctx: "Context | None" = None, bucket_dtype="fp32", algo="auto"):
        self.ctx = ctx
        self.dtype = dtype_code(bucket_dtype)
        self.algo = algo
        self._plans = {}
        self._offsets = {}
        self._next = 0
        self.launches = 0

    def slot(self, index: int, numel: int):
        ctx = self.ctx or default_context()
        self.ctx = ctx
        key = (index, numel)
        if key not in self._plans:
            plan = ctx.plan([numel], self.dtype)
            if self._next + plan.bucket_bytes > ctx.staging_bytes:
                self._next = 0
                if plan.bucket_bytes > ctx.staging_bytes:
                    raise _lib.B200DDPError("DDP bucket larger than the staging area; raise B200DDP_STAGING_MB "
                                            "or lower bucket_cap_mb")
            self._plans[key] = plan
            self._offsets[key] = self._next
            self._next = (self._next + plan.bucket_bytes + 255) // 256 * 256
        return self._plans[key], self._offsets[key]


def comm_hook(state: HookState, bucket: dist.GradBucket) -> torch.futures.Future[torch.Tensor]:
    # DDP checks this exact return annotation (torch/nn/parallel/distributed.py:2280-2290)
    buf = bucket.buffer()
    if buf.dtype != torch.float32:
        raise _lib.B200DDPError("b200ddp.comm_hook expects fp32 gradient buckets")
    plan, off = state.slot(bucket.index(), buf.numel())
    ctx = state.ctx
    n = ctx.world
    pow2 = n & (n - 1) == 0
    pre = 1.0 / n if (pow2 or state.dtype == _lib.BF16) else 1.0
    # everything is enqueued on the current stream, on which DDP consumes the result
    ctx.pack_scale(plan, [buf.data_ptr()], ctx.bucket_ptr(off), pre)
    res = ctx.allreduce(off, plan.bucket_elems, state.dtype, state.algo)
    ctx.unpack(plan, ctx.bucket_ptr(res), [buf.data_ptr()], 1.0 if pre != 1.0 or n == 1 else float(n))
    state.launches += 3 if n > 1 else 2
    fut = torch.futures.Future()
    fut.set_result(buf)
    return fut
