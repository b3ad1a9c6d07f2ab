"""CPU restatement of the reference's manual data-parallel step.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Every function cites the
reference lines it follows; `oracle/make_golden.py` checks each one against
the reference function imported from /root/reference and freezes the outputs
under tests/golden/.

The arithmetic itself lives in third-party torch (c10d all_reduce, ATen div_,
torch.optim.SGD); the reference pins torch 1.13.1
(codes/task2/docker-compose.yml:6), this image has 2.11.0 - the semantics used
here are unchanged between the two.
"""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


# --------------------------------------------------------------------------
# process group helpers: codes/task4/dist_utils.py:6-30 (the gloo twin of
# codes/task2/dist_utils.py:6-30, which hard-codes "nccl")
# --------------------------------------------------------------------------
def dist_init(world_size, rank, master_addr="127.0.0.1", master_port="12355",
              backend="gloo"):
    os.environ["MASTER_ADDR"] = str(master_addr)
    os.environ["MASTER_PORT"] = str(master_port)
    dist.init_process_group(backend=backend, rank=rank, world_size=world_size)
    assert dist.is_initialized()
    return True


def get_local_rank():
    # codes/task2/dist_utils.py:18-23 - returns the *global* rank.
    return dist.get_rank() if dist.is_initialized() else 0


def get_world_size():
    # codes/task2/dist_utils.py:26-30
    return dist.get_world_size() if dist.is_initialized() else 1


def init_parameters(model):
    # codes/task2/dist_utils.py:33-37: parameters only, buffers are not sent.
    if get_world_size() > 1:
        for param in model.parameters():
            dist.broadcast(param.data, 0)


def allreduce_average_gradients(model):
    # codes/task2/dist_utils.py:39-42: one SUM all-reduce per parameter, then
    # an in-place true division by the (int) world size.
    for param in model.parameters():
        dist.all_reduce(param.grad.data, op=dist.ReduceOp.SUM)
        param.grad.data /= get_world_size()


def average_gradients(model):
    # codes/task3/dist_utils.py:40-46: same, divisor spelled float(size).
    size = float(get_world_size())
    for param in model.parameters():
        dist.all_reduce(param.grad.data, op=dist.ReduceOp.SUM)
        param.grad.data /= size


# --------------------------------------------------------------------------
# in-process stand-in for the collective, for tests that cannot spawn:
# sum the ranks' tensors left to right (rank 0 first), then divide.
# Exact for N <= 2 (fp32 addition is commutative, there is one addition);
# for N > 2 gloo's order may differ, so callers use the stated tolerance.
# --------------------------------------------------------------------------
def divide_by_world(x, n, rule="cpu"):
    """`param.grad.data /= world_size` (codes/task2/dist_utils.py:42) on a float32 numpy array.
    rule "cpu": the correctly rounded quotient - ATen's CPU kernel, i.e. the gloo run of the lab.
    rule "cuda": x * fp32(1/n) - ATen's CUDA kernel multiplies by the reciprocal when the divisor
    is a host scalar (div_true_kernel_cuda), and the reference's scripts run there (backend "nccl",
    codes/task2/dist_utils.py:12).  Pinned on the GPU box against torch itself
    (tests/test_gpu_kernels.py::test_reciprocal_division_is_what_torch_cuda_computes).
    Both rules agree bit for bit when n is a power of two."""
    x = np.asarray(x, np.float32)
    if rule == "cuda":
        return (x * (np.float32(1.0) / np.float32(n))).astype(np.float32)
    if rule != "cpu":
        raise ValueError(rule)
    with np.errstate(all="ignore"):
        return (x / np.float32(n)).astype(np.float32)


def average_over_ranks(per_rank: list[torch.Tensor], rule="cpu") -> torch.Tensor:
    acc = per_rank[0].clone()
    for t in per_rank[1:]:
        acc += t
    if rule == "cpu":
        acc /= len(per_rank)
        return acc
    return torch.from_numpy(divide_by_world(acc.numpy(), len(per_rank), rule))


def bf16_wire_average(per_rank: list[torch.Tensor]) -> torch.Tensor:
    """torch bf16_compress_hook semantics
    (torch/distributed/algorithms/ddp_comm_hooks/default_hooks.py:57-93):
    cast to bf16, divide by N, sum in bf16, cast back to fp32."""
    n = len(per_rank)
    acc = per_rank[0].to(torch.bfloat16).div_(n)
    for t in per_rank[1:]:
        acc += t.to(torch.bfloat16).div_(n)
    return acc.to(torch.float32)


# --------------------------------------------------------------------------
# momentum SGD, restating torch/optim/sgd.py:322-380 (_single_tensor_sgd),
# which is what optimizer.step() at codes/task2/model.py:62 runs.
# add_(x, alpha=a) on CPU is a fused multiply-add (checked in make_golden.py),
# emulated here through float64 (product of two fp32 is exact in fp64).
# --------------------------------------------------------------------------
def _fma32(a, b, c):
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64)
            + np.asarray(c, np.float64)).astype(np.float32)


def sgd_step_numpy(p, g, buf, *, lr, momentum=0.0, dampening=0.0,
                   weight_decay=0.0, nesterov=False, maximize=False):
    """Returns (new_p, new_buf).  `buf is None` means first step."""
    p = np.asarray(p, np.float32)
    g = np.asarray(g, np.float32)
    if maximize:
        g = -g
    if weight_decay != 0:
        g = _fma32(np.float32(weight_decay), p, g)           # sgd.py:346-356
    if momentum != 0:
        if buf is None:
            buf = g.copy()                                    # sgd.py:361-363
        else:
            buf = (np.asarray(buf, np.float32) * np.float32(momentum)).astype(np.float32)
            buf = _fma32(np.float32(1 - dampening), g, buf)   # sgd.py:365
        g = _fma32(np.float32(momentum), buf, g) if nesterov else buf  # :367-370
    new_p = _fma32(np.float32(-lr), g, p)                     # sgd.py:380
    return new_p, buf


# --------------------------------------------------------------------------
# the lab's own optimizers, codes/task1/pytorch/MyOptimizer.py:18-43, op by op as torch
# evaluates the expressions on fp32 tensors (each tensor operation rounds once):
#   GD    p = p - lr * g                                             (:23)
#   Adam  m = b1 * m + (1 - b1) * g                                  (:39)
#         v = b2 * v + (1 - b2) * g**2                               (:41)
#         p = p - (lr / (sqrt(v) + eps)) * m                         (:42-43)
# `(1 - b1)` is evaluated by Python in double before the product casts it to fp32;
# `lr / tensor` is Tensor.__rtruediv__ = tensor.reciprocal() * lr.
# Pinned against the reference classes themselves in make_golden.py (opt_*.npz):
#   * adam_step_torch is the reference's three statements verbatim on torch tensors; on the CPU it
#     reproduces the reference class bit for bit (same ATen kernels), and on a CUDA device it is the
#     reference's arithmetic on the GPU (the GPU tests run it there).
#   * the numpy functions use IEEE sqrt / division, which is what ATen's CUDA kernels compute.
#     ATen's CPU sqrt (MKL VML) is NOT correctly rounded (0.65 % of fp32 inputs differ by one ulp,
#     measured in this container), so against the CPU goldens they are held to rel <= 1e-6 on m, v
#     and p rather than to bit equality; GD has no sqrt and is bit-exact everywhere.
# --------------------------------------------------------------------------
def adam_step_torch(p, g, m, v, *, lr, beta1=0.9, beta2=0.999, eps=1e-8):
    """codes/task1/pytorch/MyOptimizer.py:39-43 on torch tensors of any device; returns (p, m, v)."""
    m = beta1 * m + (1 - beta1) * g
    v = beta2 * v + (1 - beta2) * g ** 2
    p = p - (lr / (v.sqrt() + eps)) * m
    return p, m, v


def gd_step_torch(p, g, *, lr):
    """codes/task1/pytorch/MyOptimizer.py:23."""
    return p - lr * g


def gd_step_numpy(p, g, *, lr):
    p, g = np.asarray(p, np.float32), np.asarray(g, np.float32)
    return (p - (np.float32(lr) * g).astype(np.float32)).astype(np.float32)


def adam_step_numpy(p, g, m, v, *, lr, beta1=0.9, beta2=0.999, eps=1e-8):
    """Returns (new_p, new_m, new_v); m and v start as zeros (MyOptimizer.py:32-33)."""
    f = np.float32
    p, g, m, v = (np.asarray(a, f) for a in (p, g, m, v))
    m = ((f(beta1) * m).astype(f) + (f(1 - beta1) * g).astype(f)).astype(f)
    v = ((f(beta2) * v).astype(f) + (f(1 - beta2) * (g * g).astype(f)).astype(f)).astype(f)
    den = (np.sqrt(v).astype(f) + f(eps)).astype(f)
    step = (((f(1) / den).astype(f) * f(lr)).astype(f) * m).astype(f)
    return (p - step).astype(f), m, v


# --------------------------------------------------------------------------
# the lab CNN: same layers, shapes and parameter order as `Net` at
# codes/task2/model.py:13-37 (51,902 parameters in 8 tensors).
# --------------------------------------------------------------------------
class LabNet(nn.Module):
    def __init__(self, in_channels=1, num_classes=10):
        super().__init__()
        self.conv1 = nn.Conv2d(in_channels, 6, 5, stride=1, padding=2)
        self.conv2 = nn.Conv2d(6, 16, 5, stride=1, padding=0)
        self.fc1 = nn.Linear(16 * 5 * 5, 120)
        self.fc2 = nn.Linear(120, num_classes)

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.conv1(x)), 2)
        x = F.max_pool2d(F.relu(self.conv2(x)), 2)
        x = F.relu(self.fc1(x.flatten(1)))
        return self.fc2(x)


LABNET_NUMELS = (150, 6, 2400, 16, 48000, 120, 1200, 10)


def synthetic_batch(rank, step, batch, chw, classes):
    """Per-rank synthetic batch: ToTensor-range images
    and integer labels from a generator seeded 1234 + rank, advanced `step`
    batches."""
    g = torch.Generator().manual_seed(1234 + rank)
    for _ in range(step + 1):
        x = torch.rand((batch,) + tuple(chw), generator=g)
        y = torch.randint(0, classes, (batch,), generator=g)
    return x, y
