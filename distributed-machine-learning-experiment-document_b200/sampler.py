"""Samplers whose index shards are built on the GPU (K4), bit-exact with torch.

`DistributedSampler` stands in for torch.utils.data.DistributedSampler as the
reference uses it (codes/task2/model.py:124: dataset, num_replicas, rank;
defaults shuffle=True, seed=0, drop_last=False).  `MySampler` has the
constructor of the reference's skeleton (codes/task3/sampler.py:5-14) with
`__iter__` filled in.  The DataLoader consumes a Python iterator, so the shard
is copied to the host once per epoch (8 bytes per index).
"""
from __future__ import annotations

import torch
import torch.distributed as dist
from torch.utils.data import Dataset, Sampler

from .runtime import default_context, sampler_sizes


class DistributedSampler(Sampler):
    def __init__(self, dataset: Dataset, num_replicas=None, rank=None, shuffle=True, seed=0, drop_last=False):
        # argument handling of torch/utils/data/distributed.py:66-105
        if num_replicas is None:
            if not dist.is_available() or not dist.is_initialized():
                raise RuntimeError("Requires distributed package to be available")
            num_replicas = dist.get_world_size()
        if rank is None:
            if not dist.is_available() or not dist.is_initialized():
                raise RuntimeError("Requires distributed package to be available")
            rank = dist.get_rank()
        if rank >= num_replicas or rank < 0:
            raise ValueError(f"Invalid rank {rank}, rank should be in the interval [0, {num_replicas - 1}]")
        self.dataset = dataset
        self.num_replicas = num_replicas
        self.rank = rank
        self.epoch = 0
        self.drop_last = drop_last
        self.num_samples, self.total_size = sampler_sizes(len(dataset), num_replicas, drop_last)
        self.shuffle = shuffle
        self.seed = seed

    def device_indices(self) -> torch.Tensor:
        """This epoch's shard as an int64 CUDA tensor (no host copy)."""
        return default_context().sampler_shard(len(self.dataset), self.num_replicas, self.rank, self.seed,
                                               self.epoch, self.shuffle, self.drop_last)

    def __iter__(self):
        indices = self.device_indices().cpu().tolist()
        if indices and indices[0] < 0:
            raise RuntimeError("the device shuffle did not finish within its round budget (see include/b200ddp.h)")
        if len(indices) != self.num_samples:
            raise AssertionError(f"Number of subsampled indices ({len(indices)}) does not match "
                                 f"num_samples ({self.num_samples})")
        return iter(indices)

    def __len__(self):
        return self.num_samples

    def set_epoch(self, epoch: int):
        # torch/utils/data/distributed.py:146-157; suggested for the lab at sections/task3.tex:52
        self.epoch = epoch


class MySampler(DistributedSampler):
    """codes/task3/sampler.py:5-25 with __iter__ implemented.

    mode="partition" (default): every rank shuffles with the SAME seed and takes
    a disjoint strided shard - the lab's "random partition"
    (sections/task3.tex:21-23), identical to DistributedSampler.
    mode="sampling": every rank draws its own permutation with `seed` (the
    reference passes seed=args.rank, codes/task3/model.py:111) and keeps the
    first ceil(n / num_replicas) indices - the lab's "random sampling"; shards
    of different ranks may overlap.
    """

    def __init__(self, dataset: Dataset, num_replicas, rank, shuffle=True, seed=0, mode="partition"):
        if mode not in ("partition", "sampling"):
            raise ValueError("mode must be 'partition' or 'sampling'")
        super().__init__(dataset, num_replicas, rank, shuffle=shuffle, seed=seed)
        self.mode = mode

    def device_indices(self) -> torch.Tensor:
        if self.mode == "partition":
            return super().device_indices()
        # one-replica shard = the whole permutation; keep its head
        full = default_context().sampler_shard(len(self.dataset), 1, 0, self.seed, self.epoch, self.shuffle, False)
        return full[: self.num_samples]
