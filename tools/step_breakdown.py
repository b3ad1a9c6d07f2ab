#!/usr/bin/env python
"""Where one training step's time goes, host side and device side.

    python tools/step_breakdown.py [--model resnet18] [--steps 30] [--sync b200|torch] [--cprofile]
    python -m torch.distributed.run --nproc-per-node 2 ... tools/step_breakdown.py

Phases: forward+loss, zero_grad+backward, gradient sync + optimizer step.  For
each: host time to ENQUEUE it (perf_counter) and device time to RUN it (CUDA
events).  `--sync torch` runs the reference's own GPU path instead (one NCCL
all_reduce + div per parameter, torch.optim.SGD: codes/task2/dist_utils.py:39-42)
for a side-by-side.  Diagnostic only; bench.py is the number of record.
"""
import argparse
import cProfile
import json
import os
import pstats
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet18")
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--sync", default="b200", choices=["b200", "torch"])
    ap.add_argument("--mode", default="fused", choices=["fused", "strict"])
    ap.add_argument("--algo", default="auto")
    ap.add_argument("--host-sync", action="store_true", help="loss.item() every step, like the reference loop")
    ap.add_argument("--cprofile", action="store_true")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    import b200ddp

    world, rank, local = bench.env_int("WORLD_SIZE", 1), bench.env_int("RANK", 0), bench.env_int("LOCAL_RANK", 0)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    name, batch, lr = bench.MODELS[args.model]
    model, chw, classes = bench.build_model(name)
    model = model.to(dev).train()
    b200ddp.configure(algo=args.algo)
    if args.sync == "b200":
        opt = b200ddp.SGD(model.parameters(), lr=lr, momentum=0.9)
        if args.mode == "fused":
            b200ddp.enable_fused_step(model, opt)
        sync = lambda: b200ddp.allreduce_average_gradients(model)
    else:
        opt = torch.optim.SGD(model.parameters(), lr=lr, momentum=0.9)

        def sync():
            if world > 1:
                for p in model.parameters():
                    dist.all_reduce(p.grad.data, op=dist.ReduceOp.SUM)
                    p.grad.data /= world
    loss_fn = torch.nn.CrossEntropyLoss()
    data = [(x.to(dev), y.to(dev)) for x, y in bench.make_batches(rank, 4, batch, chw, classes, pin=False)]

    host = {"fwd": [], "bwd": [], "sync_step": [], "total": []}
    devt = {"fwd": [], "bwd": [], "sync_step": []}
    evs = []

    def ev():
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def step(i, rec):
        x, y = data[i % 4]
        t0 = time.perf_counter(); e0 = ev()
        loss = loss_fn(model(x), y)
        t1 = time.perf_counter(); e1 = ev()
        opt.zero_grad()
        loss.backward()
        t2 = time.perf_counter(); e2 = ev()
        sync()
        opt.step()
        t3 = time.perf_counter(); e3 = ev()
        if args.host_sync:
            loss.item()
        if rec:
            host["fwd"].append(t1 - t0); host["bwd"].append(t2 - t1); host["sync_step"].append(t3 - t2)
            host["total"].append(time.perf_counter() - t0)
            evs.append((e0, e1, e2, e3))

    for i in range(8):
        step(i, False)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    prof = cProfile.Profile() if args.cprofile and rank == 0 else None
    wall0 = time.perf_counter()
    if prof:
        prof.enable()
    for i in range(args.steps):
        step(i, True)
    if prof:
        prof.disable()
    torch.cuda.synchronize()
    wall = time.perf_counter() - wall0
    for e0, e1, e2, e3 in evs:
        devt["fwd"].append(e0.elapsed_time(e1) * 1e-3)
        devt["bwd"].append(e1.elapsed_time(e2) * 1e-3)
        devt["sync_step"].append(e2.elapsed_time(e3) * 1e-3)
    if rank == 0:
        out = {"model": name, "world": world, "sync": args.sync, "mode": args.mode, "host_sync": args.host_sync,
               "wall_ms_per_step": wall / args.steps * 1e3,
               "host_enqueue_ms": {k: statistics.median(v) * 1e3 for k, v in host.items()},
               "device_span_ms": {k: statistics.median(v) * 1e3 for k, v in devt.items()}}
        print(json.dumps(out), flush=True)
        if prof:
            pstats.Stats(prof).sort_stats("cumulative").print_stats(28)
    if world > 1:
        dist.barrier()
        b200ddp.reset_default_context()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
