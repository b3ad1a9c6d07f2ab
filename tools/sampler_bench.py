#!/usr/bin/env python
"""K4 timing: the on-device DistributedSampler shard vs the CPU paths.

    python tools/sampler_bench.py [--json out.json]

Per dataset length: device time of b200ddp_sampler_shard (CUDA events, median of
10 after 3 warm-ups), device time + copy of the shard to the host (what
DataLoader needs), torch.randperm and the whole torch DistributedSampler.__iter__
on one host core.  The kernel is a latency chain (MT19937
recurrence, then ~log n reservation rounds), so microseconds per epoch is the
figure of merit, not a bandwidth.
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
    ap.add_argument("--json", default="")
    args = ap.parse_args()
    import numpy as np
    import torch
    from torch.utils.data.distributed import DistributedSampler
    import b200ddp

    torch.cuda.set_device(0)
    ctx = b200ddp.Context(0, 1, 0, staging_bytes=1 << 20)
    rows = []
    for n, world in ((50000, 8), (60000, 2), (1281167, 8)):
        class DS:
            def __len__(self):
                return n
        for _ in range(3):
            ctx.sampler_shard(n, world, 0, 0, 0, True, False)
        dev, e2e = [], []
        for ep in range(10):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            a.record()
            out = ctx.sampler_shard(n, world, 0, 0, ep, True, False)
            b.record()
            host = out.cpu()
            e2e.append(time.perf_counter() - t0)
            dev.append(a.elapsed_time(b) * 1e-3)
        ref = DistributedSampler(DS(), num_replicas=world, rank=0)
        ref.set_epoch(9)
        assert np.array_equal(host.numpy(), np.asarray(list(ref), dtype=np.int64))      # bit-exact while we are here
        t_rp, t_it = [], []
        for ep in range(5):
            g = torch.Generator().manual_seed(ep)
            t0 = time.perf_counter(); torch.randperm(n, generator=g); t_rp.append(time.perf_counter() - t0)
            ref.set_epoch(ep)
            t0 = time.perf_counter(); list(ref); t_it.append(time.perf_counter() - t0)
        med = statistics.median
        rows.append({"n": n, "world": world, "device_us": med(dev) * 1e6, "device_plus_d2h_us": med(e2e) * 1e6,
                     "torch_randperm_us": med(t_rp) * 1e6, "torch_sampler_iter_us": med(t_it) * 1e6,
                     "speedup_vs_torch_randperm": med(t_rp) / med(dev),
                     "speedup_vs_torch_sampler_iter_incl_d2h": med(t_it) / med(e2e)})
        print(json.dumps(rows[-1]), flush=True)
    if args.json:
        json.dump(rows, open(args.json, "w"), indent=1)
    ctx.close()


if __name__ == "__main__":
    main()
