"""task 4's two-stage model parallelism, replicated data-parallel (BASELINE.json configs[3]).

The reference splits the lab CNN into a convolutional stage and a fully connected stage
(codes/task4/model.py:18-46), places them on two RPC workers and drives them from a third
process with distributed autograd and `DistributedOptimizer(torch.optim.SGD, ..., lr=0.01)`
(:49-66, :75-84, :126).  It has no data parallelism; `average_gradients` sits unused in
codes/task4/dist_utils.py:40-46.  Here R replicas of the two-stage pipeline run on 2 R GPUs of
one box, one process per GPU:

    rank 2k     stage 0 (conv)  --activations-->  rank 2k+1  stage 1 (fc + loss)
                                <--d(loss)/d(activations)--

and after each backward every stage averages its gradients over the R ranks that hold the same
stage - the call the reference left unused - on the B200 path: a context connected over the
stage's sub-group, pack -> all-reduce -> fused plain-SGD update (K1, K2, K3 with momentum 0).
Semantics defined here, since the reference has none for the replicated case: every replica
steps identically (replicated parameters), gradient = mean over replicas, SGD(lr) without
momentum as at :126.  With R = 1 it is the reference's computation.

Activations travel with NCCL point-to-point (torch.distributed.send/recv); the RPC framework,
its RRefs and distributed autograd are control plane and out of scope.  The batch can be cut
into micro-batches (all forwards, then all backwards, gradients accumulating in p.grad).
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

from .engine import GradSync
from .optim import SGD
from .runtime import Context


class SubNetConv(nn.Module):
    """codes/task4/model.py:18-32 (without the RRef plumbing)."""

    def __init__(self, in_channels=1):
        super().__init__()
        self.conv1 = nn.Conv2d(in_channels, 6, 5, 1, 2)
        self.conv2 = nn.Conv2d(6, 16, 5, 1, 0)

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.conv1(x)), (2, 2))
        x = F.max_pool2d(F.relu(self.conv2(x)), (2, 2))
        return x.flatten(1)


class SubNetFC(nn.Module):
    """codes/task4/model.py:34-46."""

    def __init__(self, num_classes=10):
        super().__init__()
        self.fc1 = nn.Linear(16 * 5 * 5, 120)
        self.fc2 = nn.Linear(120, num_classes)

    def forward(self, x):
        return self.fc2(F.relu(self.fc1(x)))


class TwoStageReplica:
    """One rank's share of R replicas x 2 stages.  rank = 2 * replica + stage."""

    STAGES = 2

    def __init__(self, lr=0.01, microbatches=1, in_channels=1, num_classes=10, seed=0, staging_bytes=8 << 20,
                 algo="auto"):
        if not dist.is_initialized():
            raise RuntimeError("TwoStageReplica needs an initialised process group")
        world, rank = dist.get_world_size(), dist.get_rank()
        if world % self.STAGES:
            raise ValueError("world size must be a multiple of 2 (replicas x 2 stages)")
        self.rank, self.world = rank, world
        self.replicas, self.stage, self.replica = world // self.STAGES, rank % self.STAGES, rank // self.STAGES
        self.peer = rank + 1 if self.stage == 0 else rank - 1
        self.microbatches = int(microbatches)
        self.device = torch.device("cuda", torch.cuda.current_device())
        # every rank creates every group (torch.distributed requires it); this rank keeps its stage's
        groups = [dist.new_group([r for r in range(world) if r % self.STAGES == s]) for s in range(self.STAGES)]
        self.group = groups[self.stage]
        torch.manual_seed(seed)                      # all replicas build the same weights; the whole net is built so
        conv, fc = SubNetConv(in_channels), SubNetFC(num_classes)   # that stage 1 draws the same numbers as a full model would
        self.module = (conv if self.stage == 0 else fc).to(self.device)
        self.ctx = Context(self.replica, self.replicas, self.device.index, staging_bytes=staging_bytes)
        self.ctx.connect(group=self.group)
        self.params = list(self.module.parameters())
        self.engine = GradSync(self.params, self.ctx, algo=algo)
        self.optimizer = SGD(self.params, lr=lr)      # codes/task4/model.py:126: plain SGD, no momentum
        self.optimizer._engine = self.engine
        self.optimizer._ctx = self.ctx
        self.loss_fn = nn.CrossEntropyLoss()
        self.act_shape = None

    def step(self, inputs=None, labels=None):
        """One training step.  Stage-0 ranks pass `inputs`, stage-1 ranks pass `labels` (the
        replica's batch); returns the replica's mean loss on stage-1 ranks, None on stage 0."""
        M = self.microbatches
        self.optimizer.zero_grad(set_to_none=True)
        loss_total = None
        if self.stage == 0:
            outs = []
            for x in inputs.chunk(M):
                out = self.module(x)
                dist.send(out.detach().contiguous(), self.peer)
                outs.append(out)
            for out in outs:
                g = torch.empty_like(out)
                dist.recv(g, self.peer)
                out.backward(g)
        else:
            acts, losses = [], []
            for y in labels.chunk(M):
                a = torch.empty((y.shape[0], 16 * 5 * 5), device=self.device)
                dist.recv(a, self.peer)
                a.requires_grad_(True)
                loss = self.loss_fn(self.module(a), y) / M          # mean over the whole batch
                acts.append(a)
                losses.append(loss)
            for a, loss in zip(acts, losses):
                loss.backward()
                dist.send(a.grad.contiguous(), self.peer)
            loss_total = torch.stack([l.detach() for l in losses]).sum()
        # the call the reference left unused (codes/task4/dist_utils.py:40-46), over this stage's replicas,
        # fused with the optimizer step
        if self.replicas > 1:
            self.engine.pending = True
        self.optimizer.step()
        return loss_total

    def close(self):
        self.engine.close()
        self.ctx.close()
