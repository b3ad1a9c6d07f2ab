"""GradSync: the gradient-synchronisation step over one list of parameters.

    pack (K1)  ->  all-reduce (K2)  ->  unpack            strict drop-in of
                                                          allreduce_average_gradients
    pack (K1)  ->  all-reduce (K2)  ->  unpack + SGD (K3)  fused with optimizer.step()
    pack (K1)  ->  one-shot reduce + SGD (K2+K3)           small buckets

The parameter list is cut into buckets that fit the staging area; each bucket
owns a plan (device chunk table) and a fixed staging offset.  Python only
gathers `data_ptr()`s; every byte of gradient data is moved by the kernels.
"""
from __future__ import annotations

import torch

from . import _lib
from .runtime import Context, Plan, default_context, dtype_code, sgd_args


def _is_pow2(n: int) -> bool:
    return n > 0 and (n & (n - 1)) == 0


def padded_numel(n: int) -> int:
    """Elements a tensor occupies in a bucket (offsets are 32-element aligned)."""
    return (n + _lib.BUCKET_ALIGN - 1) // _lib.BUCKET_ALIGN * _lib.BUCKET_ALIGN


def split_buckets(numels, esize: int, cap_bytes, staging_bytes: int):
    """Greedy split of a tensor list, in order, into buckets of at most cap_bytes
    that each fit the staging area.  Returns [(lo, hi, staging_offset, elems)].
    Offsets are 256-byte aligned; when the staging area is full it is reused
    from 0 (kernels of one stream run in order, and every P2P kernel ends with a
    barrier, so a bucket is never overwritten while a peer still reads it).
    Pure function of its arguments: every rank derives the same layout."""
    cap = min(cap_bytes or staging_bytes, staging_bytes)
    out, lo, elems, offset = [], 0, 0, 0

    def close(hi):
        nonlocal lo, elems, offset
        if hi == lo:
            return
        nbytes = elems * esize
        if offset + nbytes > staging_bytes:
            offset = 0
        out.append((lo, hi, offset, elems))
        offset = (offset + nbytes + 255) // 256 * 256
        lo, elems = hi, 0

    for i, n in enumerate(numels):
        need = padded_numel(n) * esize
        if need > staging_bytes:
            raise _lib.B200DDPError(f"parameter {i} ({need} bytes) does not fit the {staging_bytes}-byte "
                                    "staging area; raise B200DDP_STAGING_MB")
        if elems and (elems + padded_numel(n)) * esize > cap:
            close(i)
        elems += padded_numel(n)
    close(len(numels))
    return out


class _Bucket:
    __slots__ = ("lo", "hi", "plan", "offset", "result_offset", "algo")

    def __init__(self, lo, hi, plan, offset):
        self.lo, self.hi, self.plan, self.offset = lo, hi, plan, offset
        self.result_offset = offset
        self.algo = None


class GradSync:
    def __init__(self, params, ctx: Context | None = None, bucket_dtype="fp32", algo="auto",
                 bucket_cap_bytes: int | None = None):
        self.ctx = ctx if ctx is not None else default_context()
        self.params = [p for p in params]
        for p in self.params:
            if p.dtype != torch.float32 or not p.is_cuda or not p.is_contiguous():
                raise _lib.B200DDPError("b200ddp handles contiguous fp32 CUDA parameters "
                                        f"(got {p.dtype}, cuda={p.is_cuda}, contiguous={p.is_contiguous()})")
        self.dtype = dtype_code(bucket_dtype)
        self.algo = algo
        self.world = self.ctx.world
        esize = 2 if self.dtype == _lib.BF16 else 4
        self.buckets: list[_Bucket] = []
        for lo, hi, offset, elems in split_buckets([p.numel() for p in self.params], esize,
                                                   bucket_cap_bytes, self.ctx.staging_bytes):
            plan = self.ctx.plan([p.numel() for p in self.params[lo:hi]], self.dtype)
            assert plan.bucket_elems == elems
            self.buckets.append(_Bucket(lo, hi, plan, offset))
        self.reuses_staging = any(b.offset == 0 for b in self.buckets[1:])
        self.pending = False          # fused mode: sync requested, applied by optimizer.step()
        self.fused_optimizer = None   # weakref to the b200ddp.SGD this engine is fused with
        self.launches = 0             # kernels of ours enqueued so far

    # ---- helpers -----------------------------------------------------------------
    def _grads(self, b: _Bucket):
        out = []
        for p in self.params[b.lo:b.hi]:
            g = p.grad
            if g is None:
                # the reference has no None-guard either (codes/task2/dist_utils.py:40-42)
                raise AttributeError("'NoneType' object has no attribute 'data': a parameter has no gradient")
            if g.dtype != torch.float32 or not g.is_contiguous():
                raise _lib.B200DDPError("gradients must be contiguous fp32")
            out.append(g)
        return out

    def _scales(self):
        """(pack scale, later divisor).  Power-of-two worlds pre-scale exactly in
        the pack; others divide after the sum, as the reference does."""
        n = self.world
        if n == 1:
            return 1.0, 1.0
        if _is_pow2(n) or self.dtype == _lib.BF16:
            return 1.0 / n, 1.0
        return 1.0, float(n)

    def _pack_and_reduce(self, b: _Bucket, grad_ptrs, fuse_oneshot: bool):
        """Returns True if the reduction was left to a fused one-shot update."""
        scale, _ = self._scales()
        self.ctx.pack_scale(b.plan, grad_ptrs, self.ctx.bucket_ptr(b.offset), scale)
        self.launches += 1
        b.result_offset = b.offset
        if self.world == 1:
            return False
        algo = _lib.ALGO_NAMES[self.algo] if isinstance(self.algo, str) else int(self.algo)
        if algo == _lib.ALGO_AUTO:
            algo = self.ctx.allreduce_pick(b.plan.bucket_elems, self.dtype)
        b.algo = algo
        if fuse_oneshot and algo == _lib.ALGO_ONESHOT:
            return True
        b.result_offset = self.ctx.allreduce(b.offset, b.plan.bucket_elems, self.dtype, algo)
        self.launches += 1
        return False

    # ---- strict drop-in: p.grad <- average --------------------------------------------
    def sync(self):
        """codes/task2/dist_utils.py:39-42 for the whole parameter list."""
        if self.world == 1:
            return            # all_reduce over one rank and a division by 1 change nothing
        _, divisor = self._scales()
        for b in self.buckets:
            gs = _lib.ptr_array([g.data_ptr() for g in self._grads(b)])
            self._pack_and_reduce(b, gs, fuse_oneshot=False)
            self.ctx.unpack(b.plan, self.ctx.bucket_ptr(b.result_offset), gs, divisor)
            self.launches += 1

    # ---- fused with the optimizer ---------------------------------------------------------
    def sync_and_step(self, moms, hyper: dict, first_step: bool, write_grads: bool = True):
        """Sync fused with torch.optim.SGD.step (codes/task2/model.py:60-62): the
        averaged gradient goes from the bucket straight into the update; p.grad
        receives it too when write_grads (the state the reference leaves)."""
        _, divisor = self._scales()
        args = sgd_args(first_step=first_step, grad_divisor=divisor, **hyper)
        has_mom = hyper.get("momentum", 0.0) != 0.0
        for b in self.buckets:
            gs = _lib.ptr_array([g.data_ptr() for g in self._grads(b)])      # gathered once, used by pack and update
            fused = self._pack_and_reduce(b, gs, fuse_oneshot=True)
            ps = [p.data_ptr() for p in self.params[b.lo:b.hi]]
            ms = [m.data_ptr() for m in moms[b.lo:b.hi]] if has_mom else None
            go = gs if write_grads else None
            if fused:
                self.ctx.allreduce_sgd(b.plan, b.offset, ps, ms, go, args)
            else:
                self.ctx.unpack_sgd(b.plan, self.ctx.bucket_ptr(b.result_offset), ps, ms, go, args)
            self.launches += 1

    def close(self):
        for b in self.buckets:
            b.plan.close()
        self.buckets = []
