"""b200ddp - B200-native data-parallel gradient synchronisation.

Drop-in for the hot path of the lab scripts in
Enigmatisms/Distributed-Machine-Learning-Experiment-Document: the manual
"all-reduce every gradient, divide by world size, SGD step" sequence
(codes/task2/dist_utils.py:39-42 + codes/task2/model.py:62) and the
DistributedSampler shards (codes/task2/model.py:124), behind the same Python
names, executed by hand-written sm_100a kernels in libb200ddp.so.

Import as `b200ddp` (the directory name of this package is not a Python
identifier; the repo root carries a two-line alias package).
"""
from . import _lib, dist_utils
from ._lib import B200DDPError
from .dist_utils import (CommTimer, allgather_average_gradients, broadcast_from_rank0, allreduce_average_gradients, average_gradients, configure,
                         dist_init, enable_fused_step, get_local_rank, get_world_size, init_parameters)
from .engine import GradSync
from .graph import GraphedTrainStep
from .hooks import HookState, comm_hook
from .optim import SGD, AdamOptimizer, BaseOptimizer, GdOptimizer
from .runtime import Context, Plan, default_context, opt_args, reset_default_context, sampler_sizes, sgd_args
from .sampler import DistributedSampler, MySampler
from .task4 import SubNetConv, SubNetFC, TwoStageReplica

__all__ = ["B200DDPError", "Context", "Plan", "GradSync", "GraphedTrainStep", "SGD", "AdamOptimizer", "GdOptimizer", "BaseOptimizer", "DistributedSampler", "MySampler", "HookState",
           "comm_hook", "dist_utils", "dist_init", "get_local_rank", "get_world_size", "init_parameters",
           "allreduce_average_gradients", "average_gradients", "allgather_average_gradients", "enable_fused_step",
           "configure", "CommTimer", "broadcast_from_rank0", "default_context", "reset_default_context", "sampler_sizes", "sgd_args", "opt_args", "TwoStageReplica", "SubNetConv", "SubNetFC"]
