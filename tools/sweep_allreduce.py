#!/usr/bin/env python
"""Gradient-bucket all-reduce sweep (BASELINE.json configs[4]): 1 KiB - 1 GiB,
fp32 / bf16, P2P kernels vs the library's NCCL path vs torch's NCCL all_reduce.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tools/sweep_allreduce.py --out profiles/sweep_N.jsonl

Timing: 5 warm-ups, then `--iters` launches bracketed by CUDA events on the
launching stream; between launches the bucket is re-packed (so every launch
reduces fresh data and L2 does not hold the previous result); max over ranks.
busbw = S / t * 2 (N - 1) / N (nccl-tests convention).
"""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--min-log2", type=int, default=10)
    ap.add_argument("--max-log2", type=int, default=30)
    ap.add_argument("--step-log2", type=int, default=1)
    ap.add_argument("--oneshot-max-mb", type=int, default=64)
    ap.add_argument("--dtypes", default="fp32,bf16")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    os.environ.setdefault("B200DDP_STAGING_MB", str((1 << (args.max_log2 - 20)) + 1))
    import b200ddp

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ctx = b200ddp.default_context()
    rows = []
    for dtype in args.dtypes.split(","):
        tdt = torch.float32 if dtype == "fp32" else torch.bfloat16
        esize = 4 if dtype == "fp32" else 2
        for lg in range(args.min_log2, args.max_log2 + 1, args.step_log2):
            S = 1 << lg
            nelem = S // esize
            src = (torch.randn(nelem // (1 if dtype == "fp32" else 1), device=dev, dtype=torch.float32,
                               generator=torch.Generator(device=dev).manual_seed(rank)) * 1e-2)
            plan = ctx.plan([nelem], dtype)
            tbuf = src.to(tdt)
            algos = (["nccl", "twoshot"] + (["nvls"] if ctx.has_multicast else []) +
                     (["oneshot"] if S <= args.oneshot_max_mb << 20 else []) + ["torch_nccl"])
            for algo in algos:
                # `iters` launches back to back between two events (host launch latency is hidden
                # once the device is the bottleneck; below that the figure IS the launch rate).
                # Values grow by a factor N per in-place reduction; from 1e-2 that stays finite.
                work = tbuf.clone()
                ctx.pack_scale(plan, [src.data_ptr()], ctx.bucket_ptr(0), 1.0 / world)

                def one():
                    if algo == "torch_nccl":
                        dist.all_reduce(work)
                    else:
                        ctx.allreduce(0, plan.bucket_elems, dtype, algo)
                ts = []
                for rep in range(3):
                    for _ in range(5 if rep == 0 else 0):
                        one()
                    if algo != "torch_nccl":
                        ctx.pack_scale(plan, [src.data_ptr()], ctx.bucket_ptr(0), 1.0 / world)
                    else:
                        work.copy_(tbuf)
                    dist.barrier()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                    for _ in range(args.iters):
                        one()
                    b.record()
                    b.synchronize()
                    ts.append(a.elapsed_time(b) * 1e-3 / args.iters)
                t = torch.tensor([statistics.median(ts)], device=dev, dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                t = float(t.item())
                rows.append({"n_gpus": world, "dtype": dtype, "bytes": S, "algo": algo, "us": t * 1e6,
                             "algbw_gbs": S / t / 1e9, "busbw_gbs": S / t / 1e9 * 2 * (world - 1) / world})
                if rank == 0:
                    print(json.dumps(rows[-1]), flush=True)
            plan.close()
            del src, tbuf
            torch.cuda.empty_cache()
    if rank == 0 and args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")
    dist.barrier()
    b200ddp.reset_default_context()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
