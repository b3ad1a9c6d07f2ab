#!/usr/bin/env python
"""The lab's task2/task3 training script on the B200-native path.

Same command-line flags and the same loop as the reference scripts
(codes/task2/model-mp.py, codes/task3/model.py): a LeNet-style CNN, SGD with
momentum 0.9, batches of 32 per rank, loss printed every 20 iterations, total
communication time at the end.  What changed is exactly what INTEGRATION.md
lists - `dist_utils`, the optimizer and the sampler come from b200ddp - plus
what the environment forces: MNIST cannot be downloaded here, so the dataset
is MNIST-shaped synthetic data, and ranks are created with mp.spawn on the
GPUs of one box.

    python examples/train_lab_cnn.py --n_devices 1
    python examples/train_lab_cnn.py --n_devices 8 --graph      # loop body replayed from a CUDA graph
"""
import argparse
import os
import sys
import time

import torch
import torch.multiprocessing as mp
import torch.nn as nn
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


class Net(nn.Module):
    # layer shapes of the lab model (codes/task2/model.py:13-37)
    def __init__(self, in_channels=1, num_classes=10):
        super().__init__()
        self.conv1 = nn.Conv2d(in_channels, 6, 5, 1, 2)
        self.conv2 = nn.Conv2d(6, 16, 5, 1, 0)
        self.fc1 = nn.Linear(16 * 5 * 5, 120)
        self.fc2 = nn.Linear(120, num_classes)

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.conv1(x)), 2)
        x = F.max_pool2d(F.relu(self.conv2(x)), 2)
        return self.fc2(F.relu(self.fc1(x.flatten(1))))


class SyntheticMNIST(torch.utils.data.Dataset):
    """MNIST-shaped data with a learnable signal: the label is the brightest of ten stripes."""

    def __init__(self, n, seed=0):
        g = torch.Generator().manual_seed(seed)
        self.y = torch.randint(0, 10, (n,), generator=g)
        self.x = torch.rand(n, 1, 28, 28, generator=g) * 0.5
        for k in range(10):
            rows = slice(2 * k + 4, 2 * k + 6)
            self.x[self.y == k, :, rows, :] += 0.5

    def __len__(self):
        return len(self.y)

    def __getitem__(self, i):
        return self.x[i], self.y[i]


def train(model, dataloader, loss_fn, optimizer, args, num_epochs=2):
    import b200ddp
    from b200ddp import dist_utils
    rank = dist_utils.get_local_rank()
    print(f"Device {rank} starts training ...")
    model.train()
    dist_utils.init_parameters(model)             # codes/task2/model.py:46
    timer = b200ddp.CommTimer()
    step = b200ddp.GraphedTrainStep(model, loss_fn, optimizer) if args.graph else None
    loss_total, start = 0.0, time.time()
    for epoch in range(num_epochs):
        dataloader.sampler.set_epoch(epoch)       # suggested at sections/task3.tex:52
        for i, (inputs, labels) in enumerate(dataloader):
            if step is not None:
                loss = step(inputs.pin_memory(), labels.pin_memory())
            else:
                inputs, labels = inputs.cuda(), labels.cuda()
                loss = loss_fn(model(inputs), labels)
                optimizer.zero_grad()
                loss.backward()
                with timer:                       # codes/task2/model-mp.py:61-66, with CUDA events
                    dist_utils.allreduce_average_gradients(model)
                optimizer.step()
            loss_total += loss.item()
            if i % 20 == 19:
                print("Device: %d epoch: %d, iters: %5d, loss: %.3f" % (rank, epoch + 1, i + 1, loss_total / 20))
                loss_total = 0.0
    print("Training Finished!")
    print("Training time: {:.3f}".format(time.time() - start))
    if step is None:
        print("Total communication time: {:.6f}".format(timer.total_seconds()))


def test(model, test_loader):
    model.eval()
    correct = 0
    with torch.no_grad():
        for inputs, labels in test_loader:
            pred = model(inputs.cuda()).argmax(1)
            correct += int((pred == labels.cuda()).sum())
    n = len(test_loader.dataset)
    print("\nTest set: Accuracy: {}/{} ({:.2f}%)\n".format(correct, n, 100.0 * correct / n))
    return correct / n


def main(rank, args):
    import b200ddp
    from b200ddp import dist_utils
    if args.n_devices > 1:
        dist_utils.dist_init(args.n_devices, rank, args.master_addr, args.master_port)
    else:
        torch.cuda.set_device(int(args.gpu))
    model = Net().cuda()
    train_set, test_set = SyntheticMNIST(args.samples, seed=0), SyntheticMNIST(2000, seed=1)
    sampler = b200ddp.DistributedSampler(train_set, num_replicas=args.n_devices, rank=rank)     # built on the GPU
    train_loader = torch.utils.data.DataLoader(train_set, batch_size=32, shuffle=False, sampler=sampler, drop_last=True)
    test_loader = torch.utils.data.DataLoader(test_set, batch_size=200, shuffle=False)
    optimizer = b200ddp.SGD(model.parameters(), lr=0.01, momentum=0.9)
    if args.fused or args.graph:
        b200ddp.enable_fused_step(model, optimizer)
    train(model, train_loader, nn.CrossEntropyLoss(), optimizer, args)
    if rank == 0:
        acc = test(model, test_loader)
        assert acc > 0.5, "the model did not learn"
    if args.n_devices > 1:
        b200ddp.reset_default_context()
        torch.distributed.destroy_process_group()


def parse_args():
    p = argparse.ArgumentParser()                 # flags of codes/task2/model.py:92-102
    p.add_argument("--n_devices", default=1, type=int, help="The distributd world size.")
    p.add_argument("--rank", default=0, type=int, help="The local rank of device.")
    p.add_argument("--gpu", default="0", type=str, help="GPU ID")
    p.add_argument("--master_addr", default="127.0.0.1", type=str, help="ip of rank 0")
    p.add_argument("--master_port", default="12355", type=str, help="ip of rank 0")
    p.add_argument("--samples", default=6400, type=int, help="synthetic training-set size")
    p.add_argument("--fused", action="store_true", help="fold the sync into optimizer.step()")
    p.add_argument("--graph", action="store_true", help="replay the loop body from a CUDA graph")
    return p.parse_args()


if __name__ == "__main__":
    args = parse_args()
    if args.n_devices > 1:
        mp.spawn(main, args=(args,), nprocs=args.n_devices, join=True)      # codes/task2/model-mp.py:146-148
    else:
        main(0, args)
