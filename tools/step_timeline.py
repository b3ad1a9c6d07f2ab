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
    ap.add_argument("--pre", default="none", choices=["none", "overlap-engine", "side-stream-kernels", "side-stream-p2p"],
                    help="something to do once BEFORE the step is built (bisecting what perturbs later replays)")
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
    if args.pre != "none" and world > 1:
        ctx = b200ddp.default_context()
        numels = [1 << 20] * 12
        fake = [torch.nn.Parameter(torch.zeros(n, device=dev)) for n in numels]
        for p in fake:
            p.grad = torch.ones_like(p)
        if args.pre == "overlap-engine":
            eng = b200ddp.GradSync(fake, ctx, "fp32", "auto", overlap=True)
            eng.sync()
            torch.cuda.synchronize()
            eng.close()
        else:
            s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
            main = torch.cuda.current_stream()
            if args.pre == "side-stream-kernels":
                for s_ in (s1, s2):
                    s_.wait_stream(main)
                    with torch.cuda.stream(s_):
                        for p in fake:
                            p.grad.add_(1.0)
                    main.wait_stream(s_)
            else:
                plan = ctx.plan(numels)
                s1.wait_stream(main)
                ctx.pack_scale(plan, [p.grad.data_ptr() for p in fake], ctx.bucket_ptr(0), 1.0, s1)
                ctx.allreduce(0, plan.bucket_elems, "fp32", "twoshot", s1)
                main.wait_stream(s1)
                torch.cuda.synchronize()
                plan.close()
            torch.cuda.synchronize()
        dist.barrier()
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
        k = sorted((e for e in kern if "ts" in e), key=lambda e: e["ts"])
        n = len(k) // args.steps
        mid = k[n:2 * n] if args.steps >= 3 else k
        ours = [e for e in mid if "b200ddp" in e["name"]]
        rest = [e for e in mid if "b200ddp" not in e["name"]]
        t0 = mid[0]["ts"]
        print("SUMMARY pre=%s overlap=%s: step span %.1f us; other kernels: %d, sum %.1f us, last ends at %.1f; ours: %s" % (
            args.pre, args.overlap, mid[-1]["ts"] + mid[-1]["dur"] - t0, len(rest), sum(e["dur"] for e in rest),
            max(e["ts"] + e["dur"] for e in rest) - t0,
            [(e["name"].split("::")[-1][:18], round(e["ts"] - t0, 1), round(e["dur"], 1)) for e in ours]))
        out = {"world": world, "overlap": args.overlap, "pre": args.pre, "steps": args.steps, "kernels": kern}
        if args.out:
            json.dump(out, open(args.out, "w"))
        print("kernels recorded:", len(kern))
    if world > 1:
        dist.barrier()
        b200ddp.reset_default_context()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
