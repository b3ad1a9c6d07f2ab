#!/usr/bin/env python
"""Gradient-bucket all-reduce sweep (BASELINE.json configs[4]): 1 KiB - 1 GiB,
fp32 / bf16, P2P kernels vs the library's NCCL path vs torch's NCCL all_reduce
vs gloo on the host cores (the reference's CPU backend, codes/task4/dist_utils.py:12).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tools/sweep_allreduce.py --out profiles/sweep_N.jsonl

    --sizes 44726784,16777216   explicit bucket sizes in bytes (rounded up to 256) instead of powers of two
    --variants                  also sweep the launch settings of the NVLS / two-shot kernels
                                (loads in flight per thread, grid cap) - must run on every rank alike
    --trace                     per-block phase stamps of each algorithm at each size (see p2p.cu trace_stamp):
                                where the time of one launch goes - waiting for peers, moving data, closing barrier
    --gloo-max-mb M             gloo-on-CPU column up to M MB (0: off)

Timing: 5 warm-ups, then `--iters` launches bracketed by CUDA events on the
launching stream; max over ranks.  busbw = S / t * 2 (N - 1) / N (nccl-tests convention).
"""
import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--min-log2", type=int, default=10)
    ap.add_argument("--max-log2", type=int, default=30)
    ap.add_argument("--step-log2", type=int, default=1)
    ap.add_argument("--sizes", default="", help="comma list of bucket sizes in bytes")
    ap.add_argument("--oneshot-max-mb", type=int, default=64)
    ap.add_argument("--dtypes", default="fp32,bf16")
    ap.add_argument("--algos", default="", help="comma list; default: every algorithm available")
    ap.add_argument("--variants", action="store_true")
    ap.add_argument("--trace", action="store_true")
    ap.add_argument("--gloo-max-mb", type=int, default=0)
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    sizes = [(int(s) + 255) // 256 * 256 for s in args.sizes.split(",") if s] or \
        [1 << lg for lg in range(args.min_log2, args.max_log2 + 1, args.step_log2)]
    os.environ.setdefault("B200DDP_STAGING_MB", str((max(sizes) >> 20) + 2))
    import b200ddp

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    gloo = dist.new_group(backend="gloo") if args.gloo_max_mb > 0 else None
    ctx = b200ddp.default_context()
    if args.trace:
        ctx.set_option("trace", 1)
    rows = []

    def settings_for(algo):
        """[(label, {option: value})] - the default first."""
        out = [(algo, {})]
        if not args.variants:
            return out
        if algo == "nvls":
            for depth in (1, 2, 4):
                for blocks in (37, 74, 148, 296):
                    out.append((f"nvls/d{depth}/b{blocks}", {"nvls_depth": depth, "nvls_blocks": blocks}))
        elif algo == "twoshot":
            for blocks in (74, 148, 296, 592):
                out.append((f"twoshot/b{blocks}", {"p2p_blocks": blocks}))
        return out

    def trace_summary(nblocks):
        """Phases of the last launch on this rank, then the worst rank's numbers."""
        st = [s for s in ctx.trace_read(nblocks) if s[0] > 0 and s[3] >= s[0]]      # blocks that ran in this launch
        nblocks = len(st)
        t0 = min(s[0] for s in st)
        wait1 = [s[1] - s[0] for s in st]
        data = [s[2] - s[1] for s in st]
        wait2 = [s[3] - s[2] for s in st]
        vals = [max(s[3] for s in st) - t0, statistics.median(wait1), max(wait1), statistics.median(data), max(data),
                statistics.median(wait2), max(wait2), max(s[0] for s in st) - t0]
        t = torch.tensor(vals, device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        keys = ["span_us", "barrier1_median_us", "barrier1_max_us", "data_median_us", "data_max_us",
                "barrier2_median_us", "barrier2_max_us", "last_block_start_us"]
        return {k: round(float(v) / 1e3, 2) for k, v in zip(keys, t.tolist())} | {"blocks": nblocks}

    for dtype in args.dtypes.split(","):
        tdt = torch.float32 if dtype == "fp32" else torch.bfloat16
        esize = 4 if dtype == "fp32" else 2
        for S in sizes:
            nelem = S // esize
            src = torch.randn(nelem, device=dev, dtype=torch.float32,
                              generator=torch.Generator(device=dev).manual_seed(rank)) * 1e-2
            plan = ctx.plan([nelem], dtype)
            S = plan.bucket_bytes
            tbuf = src.to(tdt)
            algos = (["nccl", "twoshot"] + (["nvls"] if ctx.has_multicast else []) +
                     (["oneshot"] if S <= args.oneshot_max_mb << 20 else []) + ["torch_nccl"])
            if args.algos:
                algos = [a for a in algos if a in args.algos.split(",")]
            for algo in algos:
                for label, opts in settings_for(algo):
                    for k, v in opts.items():
                        ctx.set_option(k, v)
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
                    row = {"n_gpus": world, "dtype": dtype, "bytes": S, "algo": label, "us": t * 1e6,
                           "algbw_gbs": S / t / 1e9, "busbw_gbs": S / t / 1e9 * 2 * (world - 1) / world}
                    if args.trace and algo in ("twoshot", "nvls", "oneshot"):
                        nvec = S // 16 if algo == "oneshot" else -(-(S // 16) // world)
                        cap = (ctx.get_option("nvls_blocks") if algo == "nvls" else 0) or ctx.get_option("p2p_blocks") or \
                            (74 if algo == "nvls" else 2 * torch.cuda.get_device_properties(dev).multi_processor_count)
                        nblocks = max(1, min(cap, -(-nvec // 1024)))
                        torch.cuda.synchronize()
                        ctx.set_option("trace", 0).set_option("trace", 1)      # a zeroed stamp buffer
                        dist.barrier()
                        one()                                   # a launch on its own, all ranks starting together
                        row["trace"] = trace_summary(nblocks)
                    rows.append(row)
                    if rank == 0:
                        print(json.dumps(row), flush=True)
                    for k in opts:
                        ctx.set_option(k, 0)
            if gloo is not None and S <= args.gloo_max_mb << 20:
                host = tbuf.float().cpu() if dtype == "fp32" else tbuf.cpu()
                reps = 3 if S > (8 << 20) else 10
                dist.all_reduce(host.clone(), group=gloo)
                dist.barrier(group=gloo)
                t0 = time.perf_counter()
                for _ in range(reps):
                    w = host.clone()
                    dist.all_reduce(w, group=gloo)       # codes/task2/dist_utils.py:41 on the CPU backend
                    w /= world
                dist.barrier(group=gloo)
                t = torch.tensor([(time.perf_counter() - t0) / reps], dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX, group=gloo)
                t = float(t.item())
                rows.append({"n_gpus": world, "dtype": dtype, "bytes": S, "algo": "gloo_cpu", "us": t * 1e6,
                             "algbw_gbs": S / t / 1e9, "busbw_gbs": S / t / 1e9 * 2 * (world - 1) / world,
                             "cores": len(os.sched_getaffinity(0)), "includes": "clone + all_reduce + division, host tensors"})
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
