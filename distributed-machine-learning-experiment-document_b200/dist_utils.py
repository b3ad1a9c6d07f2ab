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
_CONFIG = {"bucket_dtype": os.environ.get("B200DDP_BUCKET_DTYPE", "fp32"),
           "algo": os.environ.get("B200DDP_ALGO", "auto")}


def configure(bucket_dtype=None, algo=None):
    """Wire dtype ("fp32" | "bf16") and reduction algorithm ("auto" | "nccl" |
    "oneshot" | "twoshot") used by engines created afterwards."""
    if bucket_dtype is not None:
        _CONFIG["bucket_dtype"] = bucket_dtype
    if algo is not None:
        if algo not in _lib.ALGO_NAMES:
            raise ValueError(f"algo must be one of {sorted(_lib.ALGO_NAMES)}")
        _CONFIG["algo"] = algo
    _ENGINES.clear()


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


def init_parameters(model):
    """codes/task2/dist_utils.py:33-37: rank 0's parameters to every rank
    (parameters only - buffers are not broadcast, exactly as in the reference)."""
    if get_world_size() > 1:
        for param in model.parameters():
            dist.broadcast(param.data, 0)


def engine_for(model) -> GradSync:
    eng = _ENGINES.get(model)
    if eng is None:
        eng = _ENGINES[model] = GradSync(list(model.parameters()), default_context(),
                                         bucket_dtype=_CONFIG["bucket_dtype"], algo=_CONFIG["algo"])
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


def average_gradients(model):
    """codes/task3/dist_utils.py:40-46 (same arithmetic, divisor spelled float(size))."""
    allreduce_average_gradients(model)


def allgather_average_gradients(model):
    """codes/task2/dist_utils.py:44-49 as it was meant: the mean over ranks.  The
    reference version aliases one tensor for every slot of its gather list and
    hard-codes two ranks, so it returns the last rank's gradient; that bug is
    not reproduced."""
    allreduce_average_gradients(model)


def enable_fused_step(model, optimizer, write_grads=True):
    """Fuse the sync into optimizer.step(): pack -> reduce -> unpack+SGD, the
    averaged gradient never makes a separate round trip through p.grad.
    Between allreduce_average_gradients(model) and optimizer.step() p.grad still
    holds the LOCAL gradient; after step() it holds the mean (write_grads=True)."""
    from .optim import SGD
    if not isinstance(optimizer, SGD):
        raise TypeError("enable_fused_step needs a b200ddp.SGD optimizer")
    eng = engine_for(model)
    if len(optimizer.param_groups) != 1 or [id(p) for p in optimizer.param_groups[0]["params"]] != \
            [id(p) for p in eng.params]:
        raise ValueError("fused step needs one param group holding model.parameters() in order")
    eng.fused_optimizer = weakref.ref(optimizer)
    optimizer._engine = eng
    optimizer._write_grads = bool(write_grads)
    return eng
