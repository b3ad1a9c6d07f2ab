#!/usr/bin/env python
"""Initial-weight broadcast (init_parameters, codes/task2/dist_utils.py:33-37): T NCCL broadcasts, as the
reference issues them, against the flat path (rank 0 packs | device barrier | peers unpack from rank 0's
staging over NVLink | device barrier).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29512 tools/bcast_bench.py --model resnet50 --out profiles/r02_bcast_nN.json
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
    ap.add_argument("--model", default="resnet50", choices=["labnet", "resnet18", "resnet50"])
    ap.add_argument("--out", default="")
    ap.add_argument("--iters", type=int, default=10)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import b200ddp
    import bench

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    model = bench.build_model(args.model)[0].to(dev)
    tensors = [p.data for p in model.parameters()] + [b.data for b in model.buffers()]
    flat = [t for t in tensors if t.dtype == torch.float32]

    def reference_way():
        for t in flat:
            dist.broadcast(t, 0)

    def flat_way():
        b200ddp.broadcast_from_rank0(flat)

    def timed(fn):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(args.iters):
            torch.cuda.synchronize()
            dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            b.synchronize()
            ts.append(a.elapsed_time(b) * 1e-3)
        t = torch.tensor([statistics.median(ts)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    out = {"model": args.model, "n_gpus": world, "tensors": len(flat), "bytes": sum(t.numel() for t in flat) * 4,
           "nccl_broadcast_per_tensor_us": timed(reference_way) * 1e6, "flat_p2p_us": timed(flat_way) * 1e6}
    # correctness of the timed path: rank 0's values everywhere
    with torch.no_grad():
        for i, t in enumerate(flat):
            t.fill_(float(rank * 1000 + i))
    flat_way()
    torch.cuda.synchronize()
    out["flat_ok"] = all(float(t.flatten()[0]) == float(i) for i, t in enumerate(flat) if t.numel())
    if rank == 0:
        print(json.dumps(out), flush=True)
        if args.out:
            json.dump(out, open(args.out, "w"), indent=1)
    dist.barrier()
    b200ddp.reset_default_context()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
