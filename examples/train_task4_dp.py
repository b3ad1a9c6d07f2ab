#!/usr/bin/env python
"""task 4's two-stage split of the lab CNN, replicated data-parallel on the B200 path.

The reference (codes/task4/model.py) places a convolutional stage and a fully connected stage on
two CPU RPC workers and trains ONE copy with DistributedOptimizer(torch.optim.SGD, lr=0.01).  Here
R = n_devices / 2 copies run on n_devices GPUs (rank 2k holds the conv stage, rank 2k+1 the fc stage
of replica k); after each backward every stage averages its gradients over the R ranks that hold
it - the `average_gradients` the reference ships unused (codes/task4/dist_utils.py:40-46) - through
b200ddp: sub-group context, pack -> all-reduce -> fused plain-SGD update.

    python examples/train_task4_dp.py --n_devices 4 --epochs 1

MNIST cannot be downloaded here: the data is MNIST-shaped and synthetic (label = brightest stripe).
"""
import argparse
import os
import sys

import torch
import torch.multiprocessing as mp

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def synthetic_mnist(n, seed):
    g = torch.Generator().manual_seed(seed)
    y = torch.randint(0, 10, (n,), generator=g)
    x = torch.rand(n, 1, 28, 28, generator=g) * 0.5
    for k in range(10):
        x[y == k, :, 2 * k + 4: 2 * k + 6, :] += 0.5
    return x, y


def worker(rank, args):
    import b200ddp
    b200ddp.dist_utils.dist_init(args.n_devices, rank, args.master_addr, args.master_port)
    rep = b200ddp.TwoStageReplica(lr=0.01, microbatches=args.microbatches, seed=0)
    x, y = synthetic_mnist(args.samples, seed=100 + rep.replica)       # each replica trains on its own shard
    x, y = x.cuda(), y.cuda()
    print(f"Device {rank} (replica {rep.replica}, stage {rep.stage}) starts training ...")
    for epoch in range(args.epochs):
        for i in range(0, args.samples - 31, 32):                     # batch_size=32, codes/task4/model.py:118
            loss = rep.step(inputs=x[i:i + 32] if rep.stage == 0 else None,
                            labels=y[i:i + 32] if rep.stage == 1 else None)
            if rep.stage == 1 and rep.replica == 0 and (i // 32) % 50 == 0:
                print(f"epoch {epoch} iteration {i // 32} loss {loss.item():.4f}")
    torch.cuda.synchronize()
    torch.distributed.barrier()
    print("Training Finished!") if rank == 0 else None
    rep.close()
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--n_devices", default=2, type=int, help="The distributed world size (2 x replicas).")
    ap.add_argument("--master_addr", default="127.0.0.1", type=str)
    ap.add_argument("--master_port", default="12355", type=str)
    ap.add_argument("--epochs", default=1, type=int)
    ap.add_argument("--samples", default=6400, type=int)
    ap.add_argument("--microbatches", default=4, type=int)
    a = ap.parse_args()
    mp.spawn(worker, args=(a,), nprocs=a.n_devices, join=True)
