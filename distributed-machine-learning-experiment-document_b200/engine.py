"""GradSync: the gradient-synchronisation step over one list of parameters.

    pack (K1)  ->  all-reduce (K2)  ->  unpack            strict drop-in of
                                                          allreduce_average_gradients
    pack (K1)  ->  all-reduce (K2)  ->  unpack + SGD (K3)  fused with optimizer.step()
    pack (K1)  ->  one-shot reduce + SGD (K2+K3)           small buckets

The parameter list is cut into buckets that fit the staging area; each bucket
owns a plan (device chunk table) and a fixed staging offset.  Python only
gathers `data_ptr()`s; every byte of gradient data is moved by the kernels.

overlap=True starts pack + all-reduce of a bucket from inside backward, on a
side stream, as soon as the last gradient of the bucket has been accumulated
(post-accumulate-grad hooks): buckets are cut in REVERSE parameter order, the
order in which backward produces gradients, so that only the last, small
bucket (the first layers) is still in flight when backward ends.  The
reference syncs after the whole backward (codes/task2/model.py:58-62); the
values are the same, only the time at which the work starts differs.
"""
from __future__ import annotations

import os

import torch

from . import _lib
from .runtime import Context, Plan, default_context, dtype_code, opt_args, sgd_args

# Overlap mode.  Buckets of 8 MiB: small enough that the LAST bucket (the first layers, whose reduction
# cannot hide behind anything) is a few MB, large enough that the ~10 us of barriers per cross-GPU
# kernel stay a small part of each.  Buckets that run while backward still computes use few blocks,
# like NCCL's channels: a full-GPU grid of blocks spinning on their peers takes the SMs backward needs
# (measured: 2 ranks, ResNet-18, full grids -> 2.13 ms / step, slower than no overlap at all).
DEFAULT_OVERLAP_CAP = int(float(os.environ.get("B200DDP_OVERLAP_CAP_MB", "8")) * (1 << 20))
OVERLAP_BLOCKS = int(os.environ.get("B200DDP_OVERLAP_BLOCKS", "32"))
# ... and the HBM-bound kernels of those buckets (pack, update) take at most this many blocks per SM, so that
# backward's kernels find free slots beside them (0: fill the GPU)
OVERLAP_HBM_BLOCKS_PER_SM = int(os.environ.get("B200DDP_OVERLAP_HBM_BLOCKS", "2"))


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


def split_buckets_reverse(numels, esize: int, cap_bytes, staging_bytes: int):
    """The same greedy split walked from the LAST tensor to the first: bucket 0 holds the tail of
    the list (the gradients backward produces first).  Returns [(lo, hi, staging_offset, elems)]
    in launch order; inside a bucket the tensors keep their list order."""
    n = len(numels)
    rev = split_buckets(list(reversed(numels)), esize, cap_bytes, staging_bytes)
    return [(n - hi, n - lo, off, elems) for lo, hi, off, elems in rev]


class _Bucket:
    __slots__ = ("lo", "hi", "plan", "offset", "result_offset", "algo", "ready", "launched", "fused_pending", "done")

    def __init__(self, lo, hi, plan, offset):
        self.lo, self.hi, self.plan, self.offset = lo, hi, plan, offset
        self.result_offset = offset
        self.algo = None
        self.ready = 0               # gradients of this bucket accumulated since the last sync
        self.launched = False        # pack + reduce already enqueued (overlap mode)
        self.fused_pending = False   # reduction left to the fused one-shot update
        self.done = None             # overlap: event recorded behind this bucket's reduction


class GradSync:
    def __init__(self, params, ctx: Context | None = None, bucket_dtype="fp32", algo="auto",
                 bucket_cap_bytes: int | None = None, overlap: bool = False):
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
        numels = [p.numel() for p in self.params]
        total = sum(padded_numel(n) for n in numels) * esize
        # a list so small that one fused one-shot kernel handles it gains nothing from starting early
        self.overlap = bool(overlap) and self.world > 1 and total > (1 << 20)
        if self.overlap:
            layout = split_buckets_reverse(numels, esize, bucket_cap_bytes or DEFAULT_OVERLAP_CAP, self.ctx.staging_bytes)
        else:
            layout = split_buckets(numels, esize, bucket_cap_bytes, self.ctx.staging_bytes)
        self.buckets: list[_Bucket] = []
        for lo, hi, offset, elems in layout:
            plan = self.ctx.plan(numels[lo:hi], self.dtype)
            assert plan.bucket_elems == elems
            self.buckets.append(_Bucket(lo, hi, plan, offset))
        self.reuses_staging = any(b.offset == 0 for b in self.buckets[1:])
        if self.overlap and self.reuses_staging:
            raise _lib.B200DDPError("overlap needs every bucket resident in the staging area at once; "
                                    "raise B200DDP_STAGING_MB")
        self.pending = False          # fused mode: sync requested, applied by optimizer.step()
        self.fused_optimizer = None   # weakref to the b200ddp optimizer this engine is fused with
        self.launches = 0             # kernels of ours enqueued so far
        self._hooks = []
        self._side = None
        self._next = 0                # overlap: next bucket to launch (buckets go out in index order on every rank)
        self._dirty = False           # a gradient was accumulated twice since the last sync: early results are stale
        self._p2p_blocks = 0
        if self.overlap:
            self._p2p_blocks = self.ctx.get_option("p2p_blocks")
            # two side streams: pack + reduce of bucket k+1 must not queue behind the update of bucket k
            self._side = torch.cuda.Stream(device=self.params[0].device)
            self._upd = torch.cuda.Stream(device=self.params[0].device)
            owner = {}
            for bi, b in enumerate(self.buckets):
                for p in self.params[b.lo:b.hi]:
                    owner[id(p)] = bi
            for p in self.params:
                self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(owner[id(p)])))

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

    def _pack_and_reduce(self, b: _Bucket, grad_ptrs, fuse_oneshot: bool, stream=None):
        """Returns True if the reduction was left to a fused one-shot update."""
        scale, _ = self._scales()
        self.ctx.pack_scale(b.plan, grad_ptrs, self.ctx.bucket_ptr(b.offset), scale, stream)
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
        b.result_offset = self.ctx.allreduce(b.offset, b.plan.bucket_elems, self.dtype, algo, stream)
        self.launches += 1
        return False

    # ---- overlap: buckets leave from inside backward ------------------------------------
    def _make_hook(self, bi: int):
        def hook(param):
            b = self.buckets[bi]
            b.ready += 1
            if b.ready > b.hi - b.lo:
                self._dirty = True       # second backward before a sync: what left early is stale
            elif b.ready == b.hi - b.lo:
                self._launch_ready()
        return hook

    def _launch_ready(self):
        """Enqueue pack + reduce of every bucket that is complete, strictly in bucket order (all
        ranks must issue the cross-GPU kernels in the same order), on the side stream, behind
        everything the current stream has enqueued so far."""
        main = torch.cuda.current_stream()
        forked = False
        while self._next < len(self.buckets):
            b = self.buckets[self._next]
            if b.ready < b.hi - b.lo or b.launched:
                break
            if not forked:
                self._side.wait_stream(main)
                forked = True
            gs = _lib.ptr_array([g.data_ptr() for g in self._grads(b)])
            last = self._next == len(self.buckets) - 1        # nothing left to hide behind: full grids
            with self._hidden(not last):
                b.fused_pending = self._pack_and_reduce(b, gs, fuse_oneshot=False, stream=self._side)
            b.done = torch.cuda.Event()
            b.done.record(self._side)
            b.launched = True
            self._next += 1

    def _hidden(self, on: bool):
        """Launch settings for kernels that run beside backward (see OVERLAP_* above): small grids, and
        NO programmatic dependent launch - behind a reduction that is still waiting for its peers, a
        pre-launched pack parks its blocks in griddepcontrol.wait and backward's kernels find no free
        SM slot until the reduction ends.  The last bucket (on=False) has nothing beside it and keeps
        the context's settings."""
        eng = self

        class _Ctx:
            def __enter__(self):
                eng._saved = {}
                if on:
                    eng._saved = {k: eng.ctx.get_option(k) for k in ("p2p_blocks", "hbm_blocks_per_sm", "pdl")}
                    eng.ctx.set_option("pdl", int(os.environ.get("B200DDP_OVERLAP_PDL", "0")))
                    eng.ctx.set_option("p2p_blocks", OVERLAP_BLOCKS).set_option("hbm_blocks_per_sm", OVERLAP_HBM_BLOCKS_PER_SM)

            def __exit__(self, *exc):
                for k, v in eng._saved.items():
                    eng.ctx.set_option(k, v)
                return False
        return _Ctx()

    def _finish_overlap(self):
        """After backward: whatever has not left yet leaves now (on the side stream).  Returns False
        if the early results cannot be used (a gradient changed after its bucket left) and the
        caller must run the plain path; on True the caller enqueues the per-bucket unpack / update
        on the side stream too - bucket k's update only needs bucket k's sums, so it runs while
        backward is still producing the gradients of earlier layers - and joins at the end."""
        if self._dirty:
            torch.cuda.current_stream().wait_stream(self._side)
            self._reset_overlap()
            return False
        for b in self.buckets:
            b.ready = b.hi - b.lo if all(p.grad is not None for p in self.params[b.lo:b.hi]) else b.ready
        self._launch_ready()
        if self._next < len(self.buckets):
            self._grads(self.buckets[self._next])        # raises the reference's error for a missing gradient
        return True

    def _reset_overlap(self):
        for b in self.buckets:
            b.ready, b.launched, b.fused_pending = 0, False, False
        self._next = 0
        self._dirty = False

    # ---- strict drop-in: p.grad <- average --------------------------------------------
    def sync(self):
        """codes/task2/dist_utils.py:39-42 for the whole parameter list."""
        if self.world == 1:
            return            # all_reduce over one rank and a division by 1 change nothing
        _, divisor = self._scales()
        early = self.overlap and self._finish_overlap()
        stream = self._upd if early else None
        for b in self.buckets:
            gs = _lib.ptr_array([g.data_ptr() for g in self._grads(b)])
            if not early:
                self._pack_and_reduce(b, gs, fuse_oneshot=False)
            else:
                self._upd.wait_event(b.done)
            self.ctx.unpack(b.plan, self.ctx.bucket_ptr(b.result_offset), gs, divisor, stream)
            self.launches += 1
        if early:
            torch.cuda.current_stream().wait_stream(self._upd)
        if self.overlap:
            self._reset_overlap()

    # ---- fused with the optimizer ---------------------------------------------------------
    def sync_and_step(self, moms, hyper: dict, first_step: bool, write_grads: bool = True, rule="sgd", vels=None):
        """Sync fused with the optimizer step (codes/task2/model.py:60-62): the
        averaged gradient goes from the bucket straight into the update; p.grad
        receives it too when write_grads (the state the reference leaves).
        rule "sgd": torch.optim.SGD with `hyper`; "adam" / "gd": the lab's own optimizers."""
        _, divisor = self._scales()
        if rule == "sgd":
            args = sgd_args(first_step=first_step, grad_divisor=divisor, **hyper)
            has_mom = hyper.get("momentum", 0.0) != 0.0
        else:
            args = opt_args(rule, grad_divisor=divisor, **hyper)
            has_mom = rule == "adam"
        early = self.overlap and self._finish_overlap()
        stream = self._upd if early else None
        for b in self.buckets:
            gs = _lib.ptr_array([g.data_ptr() for g in self._grads(b)])      # gathered once, used by pack and update
            fused = False
            if not early:
                fused = self._pack_and_reduce(b, gs, fuse_oneshot=(rule == "sgd"))
            else:
                self._upd.wait_event(b.done)
            ps = [p.data_ptr() for p in self.params[b.lo:b.hi]]
            ms = [m.data_ptr() for m in moms[b.lo:b.hi]] if has_mom else None
            go = gs if write_grads else None
            if early:
                with self._hidden(b is not self.buckets[-1]):
                    self._update(b, rule, ps, ms, vels, go, args, stream)
            elif fused:
                self.ctx.allreduce_sgd(b.plan, b.offset, ps, ms, go, args)
            else:
                self._update(b, rule, ps, ms, vels, go, args, stream)
            self.launches += 1
        if early:
            torch.cuda.current_stream().wait_stream(self._upd)
        if self.overlap:
            self._reset_overlap()

    def _update(self, b, rule, ps, ms, vels, go, args, stream):
        if rule != "sgd":
            vs = [v.data_ptr() for v in vels[b.lo:b.hi]] if rule == "adam" else None
            self.ctx.unpack_opt(b.plan, self.ctx.bucket_ptr(b.result_offset), ps, ms, vs, go, args, stream)
        else:
            self.ctx.unpack_sgd(b.plan, self.ctx.bucket_ptr(b.result_offset), ps, ms, go, args, stream)

    def close(self):
        for h in self._hooks:
            h.remove()
        self._hooks = []
        for b in self.buckets:
            b.plan.close()
        self.buckets = []
