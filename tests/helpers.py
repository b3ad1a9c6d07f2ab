"""Shared helpers for the GPU parity tests (everything goes through the C ABI)."""
import ctypes
from ctypes import POINTER, byref, c_size_t, c_void_p

import numpy as np
import torch

import b200ddp
from b200ddp import _lib


def dev():
    return torch.device("cuda", 0)


def to_dev(arrays):
    return [torch.from_numpy(np.ascontiguousarray(a)).to(dev()) for a in arrays]


def ptrs(tensors):
    return [t.data_ptr() for t in tensors]


def host(tensors):
    return [t.detach().cpu().numpy() for t in tensors]


class _RawDeviceArray:
    """CUDA array interface over a raw device pointer (no copy)."""

    def __init__(self, ptr, nelem, typestr):
        self.__cuda_array_interface__ = {"shape": (nelem,), "typestr": typestr, "data": (ptr, False), "version": 2}


def read_bucket(ctx, offset_bytes, nelem, dtype=torch.float32):
    """Copy nelem bucket elements out of the staging area (to the host)."""
    if nelem == 0:
        return torch.empty(0, dtype=dtype)
    torch.cuda.synchronize()
    typestr = "<f4" if dtype == torch.float32 else "<i2"
    view = torch.as_tensor(_RawDeviceArray(ctx.staging_ptr + offset_bytes, nelem, typestr),
                           device=torch.device("cuda", ctx.device))
    out = view.clone().cpu()
    return out if dtype == torch.float32 else out.view(torch.bfloat16)


def from_bucket(flat, plan):
    """Split a flat bucket (numpy) into per-tensor arrays using the plan's offsets."""
    return [flat[o:o + n] for o, n in zip(plan.offsets, plan.numels)]


class EmuWorld:
    """N ranks emulated on ONE GPU: N contexts linked through raw pointers; the
    P2P kernels of all ranks run as one cooperative launch (b200ddp_emu_*)."""

    def __init__(self, n, staging_bytes=32 << 20):
        torch.cuda.set_device(0)
        self.n = n
        self.ctxs = [b200ddp.Context(r, n, 0, staging_bytes=staging_bytes) for r in range(n)]
        for c in self.ctxs:
            c.link_local(self.ctxs)
        self._h = (c_void_p * n)(*[c.handle for c in self.ctxs])

    def allreduce(self, offset_bytes, nelem, dtype, algo):
        res = c_size_t()
        _lib.check(_lib.lib().b200ddp_emu_allreduce(self._h, self.n, int(offset_bytes), int(nelem),
                                                    b200ddp.runtime.dtype_code(dtype), _lib.ALGO_NAMES[algo],
                                                    c_void_p(torch.cuda.current_stream().cuda_stream), byref(res)),
                   "emu_allreduce")
        return int(res.value)

    def allreduce_sgd(self, plans, offset_bytes, params, moms, grads_out, args):
        n = self.n
        PP = POINTER(c_void_p)
        ph = (c_void_p * n)(*[p.handle for p in plans])
        keep = []

        def table(per_rank):
            if per_rank is None:
                return None
            rows = [_lib.ptr_array(ptrs(ts)) for ts in per_rank]
            keep.extend(rows)
            return (PP * n)(*[ctypes.cast(r, PP) for r in rows])

        _lib.check(_lib.lib().b200ddp_emu_allreduce_sgd(self._h, ph, n, int(offset_bytes), table(params), table(moms),
                                                        table(grads_out), byref(args),
                                                        c_void_p(torch.cuda.current_stream().cuda_stream)),
                   "emu_allreduce_sgd")

    def close(self):
        torch.cuda.synchronize()
        for c in self.ctxs:
            c.close()
