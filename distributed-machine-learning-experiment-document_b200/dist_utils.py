"""Drop-in for the reference's `dist_utils` module: same names, same arguments,
same observable effects (codes/task2/dist_utils.py, codes/task3/dist_utils.py).

    import b200ddp.dist_utils as dist_utils       # instead of `import dist_utils`

The training loop (codes/task2/model.py:40-72) stays as it is.
"""
from __future__ import annotations

import os
import weakref

import torch
import torch.distributed as dist

from . import _lib
from .engine import GradSync
from .runtime import default_context

_ENGINES: "weakref.WeakKeyDictionary[torch.nn.Module, GradSync]" = weakref.WeakKeyDictionary()
_GATHER_ENGINES: "weakref.WeakKeyDictionary[torch.nn.Module, GradSync]" = weakref.WeakKeyDictionary()
_CONFIG = {"bucket_dtype": os.environ.get("B200DDP_BUCKET_DTYPE", "fp32"),
           "algo": os.environ.get("B200DDP_ALGO", "auto"),
           "overlap": os.environ.get("B200DDP_OVERLAP", "0") not in ("0", "", "false"),
           "bucket_cap_bytes": int(float(os.environ["B200DDP_BUCKET_CAP_MB"]) * (1 << 20))
           if os.environ.get("B200DDP_BUCKET_CAP_MB") else None}


def configure(bucket_dtype=None, algo=None, overlap=None, bucket_cap_bytes=None):
    """Settings of the engines created afterwards: wire dtype ("fp32" | "bf16"), reduction algorithm
    ("auto" | "nccl" | "oneshot" | "twoshot" | "nvls"), and overlap.

    overlap=True starts pack + all-reduce of each bucket (reverse parameter order, bucket_cap_bytes
    each, default 25 MiB like torch DDP) from inside backward on a side stream; the call to
    allreduce_average_gradients(model) then only joins and unpacks.  Same values as the plain path
    PROVIDED p.grad is not modified between backward and the sync call (what left early would be
    stale) - the reference's loop does not (codes/task2/model.py:58-62).  Off by default."""
    if bucket_dtype is not None:
        _CONFIG["bucket_dtype"] = bucket_dtype
    if algo is not None:
        if algo not in _lib.ALGO_NAMES:
            raise ValueError(f"algo must be one of {sorted(_lib.ALGO_NAMES)}")
        _CONFIG["algo"] = algo
    if overlap is not None:
        _CONFIG["overlap"] = bool(overlap)
    if bucket_cap_bytes is not None:
        _CONFIG["bucket_cap_bytes"] = int(bucket_cap_bytes) or None
    for eng in list(_ENGINES.values()) + list(_GATHER_ENGINES.values()):
        eng.close()
    _ENGINES.clear()
    _GATHER_ENGINES.clear()


def dist_init(world_size, rank, master_addr="localhost", master_port="12355"):
    """codes/task2/dist_utils.py:6-15.  Differences forced by one-box multi-GPU:
    the process binds to GPU `rank % device_count` (the reference leaves every
    rank on GPU 0 unless CUDA_VISIBLE_DEVICES is set per process)."""
    os.environ["MASTER_ADDR"] = str(master_addr)
    os.environ["MASTER_PORT"] = str(master_port)
    if not torch.cuda.is_available():
        raise _lib.B200DDPError("b200ddp needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(rank % torch.cuda.device_count())
    dist.init_process_group(backend="nccl", rank=rank, world_size=world_size,
                            device_id=torch.device("cuda", torch.cuda.current_device()))
    assert dist.is_initialized(), "Error! The distributed env is not initialized!"
    return True


def get_local_rank():
    # codes/task2/dist_utils.py:18-23 (returns the global rank, as the reference does)
    return dist.get_rank() if dist.is_initialized() else 0


def get_world_size():
    # codes/task2/dist_utils.py:26-30
    return dist.get_world_size() if dist.is_initialized() else 1


def broadcast_from_rank0(tensors, ctx=None):
    """Rank 0's values of `tensors` to every rank, as flat buckets: rank 0 packs them into
    its staging area, the other ranks unpack straight out of rank 0's memory over NVLink
    (one pack + one unpack per bucket instead of one NCCL broadcast per tensor).
    fp32 CUDA tensors take this path; anything else (e.g. BatchNorm's int64
    num_batches_tracked) falls back to torch.distributed.broadcast."""
    from .engine import split_buckets
    ctx = ctx if ctx is not None else default_context()
    flat = [t for t in tensors if t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()]
    flat_ids = {id(t) for t in flat}
    for t in tensors:
        if id(t) not in flat_ids:
            dist.broadcast(t, dist.get_global_rank(ctx._group, 0) if ctx._group is not None else 0, group=ctx._group)
    if not flat or ctx.world == 1:
        return
    numels = [t.numel() for t in flat]
    for lo, hi, off, _ in split_buckets(numels, 4, None, ctx.staging_bytes):
        key = tuple(numels[lo:hi])
        plan = ctx._bcast_plans.get(key)
        if plan is None:
            plan = ctx._bcast_plans[key] = ctx.plan(numels[lo:hi])
        ptrs = [t.data_ptr() for t in flat[lo:hi]]
        # everything is stream-ordered; the two device-side barriers (one tiny kernel each) replace
        # the host synchronisation + process-group barriers of round 1
        if ctx.rank == 0:
            ctx.pack_scale(plan, ptrs, ctx.bucket_ptr(off), 1.0)
        ctx.barrier()                       # rank 0's bucket is complete and visible
        if ctx.rank != 0:
            ctx.unpack(plan, ctx.peer_bucket_ptr(0, off), ptrs, 1.0)
        ctx.barrier()                       # everyone has read it; the staging area may be reused


def init_parameters(model, buffers=False):
    """codes/task2/dist_utils.py:33-37: rank 0's parameters to every rank.  Like the
    reference, parameters only by default - the reference does not broadcast buffers
    (BatchNorm running statistics), which only stays harmless while every rank builds the
    model from the same seed; buffers=True sends them too."""
    if get_world_size() > 1:
        tensors = [p.data for p in model.parameters()]
        if buffers:
            tensors += [b.data for b in model.buffers()]
        broadcast_from_rank0(tensors)


def engine_for(model) -> GradSync:
    eng = _ENGINES.get(model)
    if eng is None:
        eng = _ENGINES[model] = GradSync(list(model.parameters()), default_context(),
                                         bucket_dtype=_CONFIG["bucket_dtype"], algo=_CONFIG["algo"],
                                         overlap=_CONFIG["overlap"], bucket_cap_bytes=_CONFIG["bucket_cap_bytes"])
    return eng


def allreduce_average_gradients(model):
    """codes/task2/dist_utils.py:39-42: afterwards every p.grad holds the mean
    over ranks.  With enable_fused_step(model, optimizer) the work is deferred
    to optimizer.step(), which then also leaves the mean in p.grad."""
    eng = engine_for(model)
    if eng.fused_optimizer is not None and eng.fused_optimizer() is None:
        eng.fused_optimizer = None        # the optimizer it was fused with is gone: back to the strict path
    if eng.fused_optimizer is not None:
        for p in eng.params:
            if p.grad is None:
                raise AttributeError("'NoneType' object has no attribute 'data'")
        eng.pending = True
        return
    eng.sync()
    _straggle()


def _straggle():
    """The lab's bottleneck-node experiment (sections/task2.tex:19; the commented-out
    time.sleep(0.1) on rank 1 at codes/task2/model-mp.py:47,64-65):
    B200DDP_STRAGGLER_RANK=r B200DDP_STRAGGLER_MS=t delays rank r by t ms after each sync."""
    r = os.environ.get("B200DDP_STRAGGLER_RANK")
    if r is not None and int(r) == get_local_rank():
        import time
        time.sleep(float(os.environ.get("B200DDP_STRAGGLER_MS", "100")) / 1e3)


def average_gradients(model):
    """codes/task3/dist_utils.py:40-46 (same arithmetic, divisor spelled float(size))."""
    allreduce_average_gradients(model)


def allgather_average_gradients(model):
    """codes/task2/dist_utils.py:44-49 as it was meant: every rank gathers every rank's
    gradient and averages locally.  On NVLink that is the one-shot kernel (each GPU reads all
    N buckets, (N-1) x S inbound) - the all-reduce variant above moves 2 (N-1)/N x S - so the
    lab's AllReduce-vs-AllGather cost comparison (sections/checking.tex:20-22) is
    allreduce_average_gradients vs this.  The reference version aliases one tensor for every
    slot of its gather list and hard-codes two ranks, so it returns the last rank's gradient;
    that bug is not reproduced."""
    eng = _GATHER_ENGINES.get(model)
    if eng is None:
        eng = _GATHER_ENGINES[model] = GradSync(list(model.parameters()), default_context(),
                                                bucket_dtype=_CONFIG["bucket_dtype"], algo="oneshot")
    eng.sync()
    _straggle()


class CommTimer:
    """Accumulated device time of the gradient sync, measured with CUDA events on the
    current stream as the handbook recommends (sections/task2.tex:71-80), in place of the
    wall-clock timer of codes/task2/model-mp.py:61-66,79 (which only times the enqueue).

        timer = b200ddp.CommTimer()
        with timer: dist_utils.allreduce_average_gradients(model)
        print("Total communication time:", timer.total_seconds())
    """

    def __init__(self):
        self._pairs = []

    def __enter__(self):
        self._a = torch.cuda.Event(enable_timing=True)
        self._a.record()
        return self

    def __exit__(self, *exc):
        b = torch.cuda.Event(enable_timing=True)
        b.record()
        self._pairs.append((self._a, b))
        return False

    def total_seconds(self) -> float:
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in self._pairs) * 1e-3


def enable_fused_step(model, optimizer, write_grads=True):
    """Fuse the sync into optimizer.step(): pack -> reduce -> unpack+SGD, the
    averaged gradient never makes a separate round trip through p.grad.
    Between allreduce_average_gradients(model) and optimizer.step() p.grad still
    holds the LOCAL gradient; after step() it holds the mean (write_grads=True)."""
    from .optim import SGD, BaseOptimizer
    if not isinstance(optimizer, (SGD, BaseOptimizer)):
        raise TypeError("enable_fused_step needs a b200ddp optimizer (SGD, AdamOptimizer, GdOptimizer)")
    eng = engine_for(model)
    held = optimizer.params if isinstance(optimizer, BaseOptimizer) else \
        (optimizer.param_groups[0]["params"] if len(optimizer.param_groups) == 1 else [])
    if [id(p) for p in held] != [id(p) for p in eng.params]:
        raise ValueError("fused step needs one param group holding model.parameters() in order")
    eng.fused_optimizer = weakref.ref(optimizer)
    optimizer._engine = eng
    optimizer._write_grads = bool(write_grads)
    return eng
