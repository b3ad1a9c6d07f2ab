#!/usr/bin/env python
"""Kernel timeline of a few replayed training steps on rank 0 (torch profiler / CUPTI activity
records; not ncu - nothing is replayed).  Shows which kernels of which stream run when, i.e.
whether the bucket reductions really overlap backward.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29547 tools/step_timeline.py --overlap on --out gpurun_out/timeline.json
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--overlap", default="on", choices=["on", "off"])
    ap.add_argument("--model", default="resnet18")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import b200ddp
    import bench

    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    name, batch, lr = bench.MODELS[args.model]
    b200ddp.configure(bucket_dtype="bf16" if args.model == "resnet50" else "fp32", overlap=(args.overlap == "on" and world > 1))
    model, chw, classes = bench.build_model(name)
    model = model.to(dev).train()
    b200ddp.init_parameters(model)
    opt = b200ddp.SGD(model.parameters(), lr=lr, momentum=0.9)
    b200ddp.enable_fused_step(model, opt)
    loss_fn = torch.nn.CrossEntropyLoss()
    step = b200ddp.GraphedTrainStep(model, loss_fn, opt)
    data = [(x.to(dev), y.to(dev)) for x, y in bench.make_batches(rank, 4, batch, chw, classes, pin=False)]
    for i in range(12):
        step(*data[i % 4])
    torch.cuda.synchronize()
    assert step.graph is not None, step.capture_error
    if world > 1:
        dist.barrier()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for i in range(args.steps):
            step(*data[i % 4])
        torch.cuda.synchronize()
    if rank == 0:
        evs = []
        for e in prof.events():
            if e.device_type == torch.autograd.DeviceType.CUDA:
                evs.append({"name": e.name[:90], "start_us": e.time_range.start, "dur_us": e.time_range.elapsed_us(),
                            "stream": getattr(e, "device_index", None)})
        try:
            path = (args.out or "timeline.json").replace(".json", "_chrome.json")
            prof.export_chrome_trace(path)
            raw = json.load(open(path))
            kern = [{"name": t["name"][:90], "ts": t["ts"], "dur": t["dur"], "stream": t.get("args", {}).get("stream")}
                    for t in raw.get("traceEvents", []) if t.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
            os.remove(path)
        except Exception as ex:
            kern = [{"error": str(ex)}]
        out = {"world": world, "overlap": args.overlap, "steps": args.steps, "kernels": kern}
        if args.out:
            json.dump(out, open(args.out, "w"))
        print("kernels recorded:", len(kern))
    if world > 1:
        dist.barrier()
        b200ddp.reset_default_context()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
