#!/usr/bin/env python
"""pack / unpack / update kernels alone on a model-shaped tensor list (one GPU).

    python tools/kernel_bench.py [--model resnet18] [--json out.json] [--only update_direct,...]

Timing method: bench.KernelTimer (three rotating operand sets, back-to-back
launches between CUDA events).  Also the command the ncu captures in profiles/
are taken with (`-k regex:update_kernel`, `-k regex:pack_kernel`).  torch's own
foreach-SGD and a plain contiguous copy of the same bytes are timed beside, the
same way.
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
    ap.add_argument("--model", default="resnet18", choices=["labnet", "resnet18", "resnet50", "flat256m"])
    ap.add_argument("--json", default="")
    ap.add_argument("--only", default="", help="comma list of kernel names to run")
    ap.add_argument("--no-torch", action="store_true")
    ap.add_argument("--variants", action="store_true",
                    help="also sweep the launch settings: copy_unroll 2/4/8 x pdl 0/1 for pack/unpack, pdl 0/1 for the update")
    args = ap.parse_args()
    import torch
    import b200ddp
    import bench

    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    numels = [64 << 20] if args.model == "flat256m" else [p.numel() for p in bench.build_model(args.model)[0].parameters()]
    P = sum(numels)
    peak = 6549.4
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except (OSError, ValueError, KeyError):
        pass
    ctx = b200ddp.Context(0, 1, 0, staging_bytes=max(64 << 20, 3 * (4 * P + (1 << 20))))
    kt = bench.KernelTimer(ctx, numels, dev, peak)
    only = [s for s in args.only.split(",") if s]
    out = {"model": args.model, "P": P, "tensors": len(numels), "peak_gbs": peak,
           "settings": {k: ctx.get_option(k) for k in ("copy_unroll", "pdl", "div_mode")},
           "kernels": kt.all(only)}
    if args.variants:
        out["variants"] = []
        default = {k: ctx.get_option(k) for k in ("copy_unroll", "pdl")}
        for pdl in (0, 1):
            for unroll in (2, 4, 8):
                ctx.set_option("pdl", pdl).set_option("copy_unroll", unroll)
                kt.close()                      # plans carry the copy chunk size: rebuild them
                kt.plans = {}
                names = ["pack_fp32", "pack_bf16", "unpack_fp32"] + (["update_direct", "update_from_bucket_write_grad"]
                                                                    if unroll == 4 else [])
                row = {"pdl": pdl, "copy_unroll": unroll, "kernels": kt.all(names)}
                out["variants"].append(row)
                print("variant", json.dumps(row), flush=True)
        ctx.set_option("pdl", default["pdl"]).set_option("copy_unroll", default["copy_unroll"])
        kt.close()
        kt.plans = {}
    for k, v in out["kernels"].items():
        print(k, v, flush=True)
    if not args.no_torch and not only:
        opts = []
        for k in range(kt.SETS):
            params = [torch.nn.Parameter(t) for t in kt.p[k]]
            for q, gr in zip(params, kt.g[k]):
                q.grad = gr
            o = torch.optim.SGD(params, lr=0.01, momentum=0.9)
            o.step()
            opts.append(o)
        src = [torch.empty(P, device=dev) for _ in range(kt.SETS)]
        dst = [torch.empty(P, device=dev) for _ in range(kt.SETS)]
        out["kernels"]["torch_foreach_sgd"] = kt.run(lambda k: opts[k].step(), 20 * P)
        out["kernels"]["torch_contiguous_copy"] = kt.run(lambda k: dst[k].copy_(src[k]), 8 * P)
        for k in ("torch_foreach_sgd", "torch_contiguous_copy"):
            print(k, out["kernels"][k], flush=True)
    if args.json:
        json.dump(out, open(args.json, "w"), indent=1)
    torch.cuda.synchronize()
    kt.close()
    ctx.close()


if __name__ == "__main__":
    main()
