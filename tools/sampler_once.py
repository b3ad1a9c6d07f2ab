#!/usr/bin/env python
"""One sampler shard per n (profiling helper: `ncu -k regex:sampler_` on this)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, b200ddp
torch.cuda.set_device(0)
ctx = b200ddp.Context(0, 1, 0, staging_bytes=1 << 20)
for n in [int(a) for a in sys.argv[1:]] or [60000]:
    for ep in range(3):
        out = ctx.sampler_shard(n, 1, 0, 0, ep, True, False)
    torch.cuda.synchronize()
    print(n, out[:4].tolist())
ctx.close()
