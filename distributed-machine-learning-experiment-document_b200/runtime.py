"""Context and plan objects over the C ABI, and the per-process default context.

One process per GPU: the context is created on the current CUDA device with
the rank / world size of the default torch.distributed group (1 rank when no
group is initialised, as the reference's get_world_size() does -
codes/task2/dist_utils.py:26-30).  Peer-memory handles and the NCCL id travel
through torch.distributed; nothing else does.
"""
from __future__ import annotations

import ctypes
import os
import weakref
from ctypes import byref, c_int64, c_size_t, c_void_p

import torch
import torch.distributed as dist

from . import _lib
from ._lib import B200DDPError, OptArgs, SgdArgs, check, ptr_array

_DTYPES = {"fp32": _lib.F32, "float32": _lib.F32, torch.float32: _lib.F32,
           "bf16": _lib.BF16, "bfloat16": _lib.BF16, torch.bfloat16: _lib.BF16,
           _lib.F32: _lib.F32, _lib.BF16: _lib.BF16}
_ESIZE = {_lib.F32: 4, _lib.BF16: 2}


def dtype_code(d) -> int:
    try:
        return _DTYPES[d]
    except KeyError:
        raise ValueError(f"bucket dtype must be fp32 or bf16, got {d!r}") from None


def nccl_mode() -> str:
    """B200DDP_NCCL: "lazy" (default), "1" (join at connect), "0" (never)."""
    m = os.environ.get("B200DDP_NCCL", "lazy")
    return m if m in ("lazy", "1", "0") else "lazy"


def _require_cuda():
    if not torch.cuda.is_available():
        raise B200DDPError("b200ddp needs a CUDA device (sm_100a); there is no CPU fallback")


def _stream_handle(stream=None) -> c_void_p:
    s = stream if stream is not None else torch.cuda.current_stream()
    return c_void_p(s.cuda_stream)


class Plan:
    """Bucket layout of a list of fp32 tensors (immutable)."""

    def __init__(self, ctx: "Context", numels, bucket_dtype="fp32"):
        self.ctx = ctx
        self.numels = [int(n) for n in numels]
        self.T = len(self.numels)
        self.dtype = dtype_code(bucket_dtype)
        self.esize = _ESIZE[self.dtype]
        arr = (c_int64 * max(self.T, 1))(*self.numels)
        h = c_void_p()
        check(_lib.lib().b200ddp_plan_create(ctx.handle, self.T, arr, self.dtype, byref(h)), "plan_create")
        self.handle = h
        self.bucket_elems = int(_lib.lib().b200ddp_plan_bucket_elems(h))
        offs = (c_int64 * max(self.T, 1))()
        check(_lib.lib().b200ddp_plan_offsets(h, offs), "plan_offsets")
        self.offsets = list(offs)[: self.T]
        self.bucket_bytes = self.bucket_elems * self.esize
        ctx._plans.add(self)

    def close(self):
        if getattr(self, "handle", None):
            _lib.lib().b200ddp_plan_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Context:
    """Per-GPU state: peer-visible staging, barrier flags, NCCL communicator."""

    def __init__(self, rank=0, world=1, device=None, staging_bytes=None):
        _require_cuda()
        if device is None:
            device = torch.cuda.current_device()
        if staging_bytes is None:
            staging_bytes = int(os.environ.get("B200DDP_STAGING_MB", "256")) << 20
        self.rank, self.world, self.device = int(rank), int(world), int(device)
        h = c_void_p()
        check(_lib.lib().b200ddp_create(self.rank, self.world, self.device, int(staging_bytes), byref(h)), "create")
        self.handle = h
        p, n = c_void_p(), c_size_t()
        check(_lib.lib().b200ddp_staging(h, byref(p), byref(n)), "staging")
        self.staging_ptr, self.staging_bytes = int(p.value), int(n.value)
        f = c_void_p()
        check(_lib.lib().b200ddp_flags(h, byref(f)), "flags")
        self.flags_ptr = int(f.value)
        self.has_nccl = False
        self.has_peers = self.world == 1
        self.has_multicast = False
        self.symmetric_error = None
        self._symmetric = None                 # (tensor, handle) of adopted torch symmetric memory
        self._group = None                     # process group the peers were connected over (None: default group)
        self._ipc = False                      # peers opened over CUDA IPC (teardown is collective then)
        self._plans = weakref.WeakSet()
        self._bcast_plans = {}                 # flat-broadcast plans, by tensor-size list

    # ---- set-up --------------------------------------------------------------
    def connect(self, group=None, nccl=True, p2p=True):
        """Collective over `group`: open every peer's staging (CUDA IPC) and join
        the library's own NCCL communicator."""
        if self.world == 1:
            return self
        self._group = group
        if p2p and not self.has_peers:
            mine = ctypes.create_string_buffer(_lib.IPC_BYTES)
            check(_lib.lib().b200ddp_ipc_export(self.handle, mine), "ipc_export")
            handles = [None] * self.world
            dist.all_gather_object(handles, bytes(mine.raw), group=group)
            blob = ctypes.create_string_buffer(b"".join(handles), _lib.IPC_BYTES * self.world)
            check(_lib.lib().b200ddp_ipc_open_peers(self.handle, blob), "ipc_open_peers")
            self.has_peers = True
            self._ipc = True
        # The library's own NCCL communicator is joined on first use of the NCCL algorithm, not here:
        # measured on 8 x B200 (profiles/r02_second_nccl_comm_n8.md), a SECOND NVLS-capable communicator in
        # the process (torch's process group holds the first) puts ~0.35 us of idle time between consecutive
        # kernels of every later CUDA-graph replay - 125 us on the 320 kernels of a ResNet-18 step.
        # B200DDP_NCCL=1 joins now (then AUTO may pick NCCL for fp32 buckets above 128 MB on more than
        # 4 GPUs without a multicast mapping); B200DDP_NCCL=0 never joins.
        if nccl and nccl_mode() == "1":
            self._connect_nccl()
        if p2p and os.environ.get("B200DDP_SYMMETRIC", "1") != "0":
            self._adopt_symmetric(group)
        return self

    def _connect_nccl(self):
        """Join the library's own NCCL communicator (collective over the context's group)."""
        if self.has_nccl or self.world == 1:
            return
        group = self._group
        uid = ctypes.create_string_buffer(_lib.NCCL_UID_BYTES)
        if self.rank == 0:
            check(_lib.lib().b200ddp_nccl_unique_id(uid), "nccl_unique_id")
        box = [bytes(uid.raw)]
        dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        uid = ctypes.create_string_buffer(box[0], _lib.NCCL_UID_BYTES)
        check(_lib.lib().b200ddp_nccl_init(self.handle, uid), "nccl_init")
        self.has_nccl = True

    def _adopt_symmetric(self, group):
        """Move the staging area into torch symmetric memory, which also yields a multicast
        address over all ranks' buffers (NVLink SHARP) for the in-switch all-reduce.  Every step
        is agreed on by all ranks; if anything is missing the CUDA-IPC staging stays in use."""
        grp = group if group is not None else dist.group.WORLD

        def all_ok(flag):
            votes = [None] * self.world
            dist.all_gather_object(votes, bool(flag), group=group)
            return all(votes)

        tensor = handle = None
        nbytes = 2 * self.staging_bytes + 256
        try:
            import torch.distributed._symmetric_memory as symm
            tensor = symm.empty(nbytes // 4, dtype=torch.float32, device=torch.device("cuda", self.device))
        except Exception as e:          # old torch, unsupported driver, ...
            self.symmetric_error = e
        if not all_ok(tensor is not None):
            return
        try:
            handle = symm.rendezvous(tensor, grp)
            tensor.zero_()
            torch.cuda.synchronize()
            ptrs = [int(p) for p in handle.buffer_ptrs]
            assert len(ptrs) == self.world
            shift = tensor.data_ptr() - ptrs[self.rank]      # the tensor may sit at an offset inside the block
            assert 0 <= shift and shift % 256 == 0
            ptrs = [p + shift for p in ptrs]
            mc = int(getattr(handle, "multicast_ptr", 0) or 0)
            mc = mc + shift if mc else 0
        except Exception as e:
            self.symmetric_error = e
            handle = None
        if not all_ok(handle is not None):
            return
        check(_lib.lib().b200ddp_adopt_symmetric(self.handle, ptr_array(ptrs), c_void_p(mc) if mc else None, nbytes),
              "adopt_symmetric")
        dist.barrier(group=group)       # everyone zeroed, adopted and closed its mappings of the old areas ...
        check(_lib.lib().b200ddp_free_retired(self.handle), "free_retired")     # ... so the old area can go
        self._symmetric = (tensor, handle)
        self.staging_ptr = tensor.data_ptr()
        self.has_multicast = bool(mc)

    # ---- settings ------------------------------------------------------------
    def set_option(self, name: str, value) -> "Context":
        """Per-context settings (include/b200ddp.h, B200DDP_OPT_*): div_mode ("ieee" | "reciprocal"), pdl,
        nvls_min_bytes, oneshot_max_bytes, peer_timeout_ms, p2p_blocks, nvls_depth, copy_unroll, trace.
        Settings that shape a collective launch must be set to the same value on every rank."""
        if name == "div_mode" and isinstance(value, str):
            value = {"ieee": _lib.DIV_IEEE, "cpu": _lib.DIV_IEEE, "reciprocal": _lib.DIV_RECIPROCAL,
                     "cuda": _lib.DIV_RECIPROCAL}[value]
        check(_lib.lib().b200ddp_set_option(self.handle, _lib.OPTIONS[name], int(value)), f"set_option({name})")
        return self

    def get_option(self, name: str) -> int:
        v = ctypes.c_longlong()
        check(_lib.lib().b200ddp_get_option(self.handle, _lib.OPTIONS[name], byref(v)), f"get_option({name})")
        return int(v.value)

    def trace_read(self, nblocks: int):
        """[(entry, barrier1, data, barrier2)] in ns for the first nblocks blocks of the last cross-GPU launch."""
        buf = (ctypes.c_uint64 * (4 * nblocks))()
        check(_lib.lib().b200ddp_trace_read(self.handle, buf, int(nblocks)), "trace_read")
        return [tuple(buf[4 * i: 4 * i + 4]) for i in range(nblocks)]

    def link_local(self, peers: "list[Context]"):
        """Same-process peers (one process driving several contexts)."""
        st = ptr_array([c.staging_ptr for c in peers])
        fl = ptr_array([c.flags_ptr for c in peers])
        check(_lib.lib().b200ddp_set_peers_direct(self.handle, st, fl), "set_peers_direct")
        self.has_peers = True
        return self

    def close(self, collective=True):
        """Frees the context.  With peers opened over CUDA IPC the teardown is collective: every rank
        closes its mappings of the others' memory, all ranks meet at a barrier, then each frees its
        own (an exported area must outlive its importers).  collective=False (garbage collection,
        interpreter exit) skips the barrier."""
        if getattr(self, "handle", None):
            for pl in list(self._plans):
                pl.close()
            if collective and self.world > 1 and self._ipc and self._group_alive():
                try:
                    check(_lib.lib().b200ddp_close_peers(self.handle), "close_peers")
                    dist.barrier(group=self._group)
                except Exception:
                    pass
            _lib.lib().b200ddp_destroy(self.handle)
            self.handle = None
            self._symmetric = None

    def _group_alive(self):
        try:
            return dist.is_available() and dist.is_initialized() and (self._group is None or dist.get_rank(self._group) >= 0)
        except Exception:
            return False

    def __del__(self):
        try:
            self.close(collective=False)
        except Exception:
            pass

    # ---- views ---------------------------------------------------------------
    def plan(self, numels, bucket_dtype="fp32") -> Plan:
        return Plan(self, numels, bucket_dtype)

    def bucket_ptr(self, offset_bytes: int) -> int:
        return self.staging_ptr + int(offset_bytes)

    def peer_bucket_ptr(self, peer: int, offset_bytes: int) -> int:
        """Address, in this process, of rank `peer`'s staging area + offset."""
        p = c_void_p()
        check(_lib.lib().b200ddp_peer_staging(self.handle, int(peer), byref(p)), "peer_staging")
        return int(p.value) + int(offset_bytes)

    # ---- kernels ---------------------------------------------------------------
    def pack_scale(self, plan: Plan, grad_ptrs, bucket_ptr: int, scale: float, stream=None):
        check(_lib.lib().b200ddp_pack_scale(self.handle, plan.handle, ptr_array(grad_ptrs), c_void_p(bucket_ptr),
                                            float(scale), _stream_handle(stream)), "pack_scale")

    def unpack(self, plan: Plan, bucket_ptr: int, grad_out_ptrs, divisor: float = 1.0, stream=None):
        check(_lib.lib().b200ddp_unpack(self.handle, plan.handle, c_void_p(bucket_ptr), ptr_array(grad_out_ptrs),
                                        float(divisor), _stream_handle(stream)), "unpack")

    def allreduce(self, offset_bytes: int, nelem: int, dtype, algo="auto", stream=None) -> int:
        """Returns the staging offset (bytes) at which the sums are."""
        res = c_size_t()
        a = _lib.ALGO_NAMES[algo] if isinstance(algo, str) else int(algo)
        if a == _lib.ALGO_NCCL and not self.has_nccl and nccl_mode() == "lazy":
            self._connect_nccl()            # every rank asks for the NCCL algorithm at the same call: collective
        check(_lib.lib().b200ddp_allreduce(self.handle, int(offset_bytes), int(nelem), dtype_code(dtype), a,
                                           _stream_handle(stream), byref(res)), "allreduce")
        return int(res.value)

    def barrier(self, stream=None):
        """Device-side barrier over all ranks on `stream` (no host synchronisation)."""
        check(_lib.lib().b200ddp_barrier(self.handle, _stream_handle(stream)), "barrier")

    def allreduce_pick(self, nelem: int, dtype) -> int:
        return int(_lib.lib().b200ddp_allreduce_pick(self.handle, int(nelem), dtype_code(dtype)))

    def unpack_sgd(self, plan: Plan, bucket_ptr: int, param_ptrs, mom_ptrs, grad_out_ptrs, args: SgdArgs, stream=None):
        check(_lib.lib().b200ddp_unpack_sgd(
            self.handle, plan.handle, c_void_p(bucket_ptr), ptr_array(param_ptrs),
            ptr_array(mom_ptrs) if mom_ptrs is not None else None,
            ptr_array(grad_out_ptrs) if grad_out_ptrs is not None else None,
            byref(args), _stream_handle(stream)), "unpack_sgd")

    def sgd_step(self, plan: Plan, grad_ptrs, param_ptrs, mom_ptrs, args: SgdArgs, stream=None):
        check(_lib.lib().b200ddp_sgd_step(
            self.handle, plan.handle, ptr_array(grad_ptrs), ptr_array(param_ptrs),
            ptr_array(mom_ptrs) if mom_ptrs is not None else None, byref(args), _stream_handle(stream)), "sgd_step")

    def unpack_opt(self, plan: Plan, bucket_ptr: int, param_ptrs, state1_ptrs, state2_ptrs, grad_out_ptrs,
                   args: OptArgs, stream=None):
        check(_lib.lib().b200ddp_unpack_opt(
            self.handle, plan.handle, c_void_p(bucket_ptr), ptr_array(param_ptrs),
            ptr_array(state1_ptrs) if state1_ptrs is not None else None,
            ptr_array(state2_ptrs) if state2_ptrs is not None else None,
            ptr_array(grad_out_ptrs) if grad_out_ptrs is not None else None,
            byref(args), _stream_handle(stream)), "unpack_opt")

    def opt_step(self, plan: Plan, grad_ptrs, param_ptrs, state1_ptrs, state2_ptrs, args: OptArgs, stream=None):
        check(_lib.lib().b200ddp_opt_step(
            self.handle, plan.handle, ptr_array(grad_ptrs), ptr_array(param_ptrs),
            ptr_array(state1_ptrs) if state1_ptrs is not None else None,
            ptr_array(state2_ptrs) if state2_ptrs is not None else None,
            byref(args), _stream_handle(stream)), "opt_step")

    def allreduce_sgd(self, plan: Plan, offset_bytes: int, param_ptrs, mom_ptrs, grad_out_ptrs, args: SgdArgs,
                      stream=None):
        check(_lib.lib().b200ddp_allreduce_sgd(
            self.handle, plan.handle, int(offset_bytes), ptr_array(param_ptrs),
            ptr_array(mom_ptrs) if mom_ptrs is not None else None,
            ptr_array(grad_out_ptrs) if grad_out_ptrs is not None else None,
            byref(args), _stream_handle(stream)), "allreduce_sgd")

    def sampler_shard(self, n, world, rank, seed, epoch, shuffle, drop_last, stream=None) -> torch.Tensor:
        """int64 CUDA tensor holding this rank's indices."""
        num = sampler_sizes(n, world, drop_last)[0]
        # rank / geometry errors must surface even when the shard is empty
        out = torch.empty(max(num, 0), dtype=torch.int64, device=torch.device("cuda", self.device))
        check(_lib.lib().b200ddp_sampler_shard(self.handle, int(n), int(world), int(rank),
                                               int(seed) & 0xFFFFFFFFFFFFFFFF, int(epoch) & 0xFFFFFFFFFFFFFFFF,
                                               int(bool(shuffle)), int(bool(drop_last)),
                                               c_void_p(out.data_ptr()) if num > 0 else None,
                                               _stream_handle(stream)), "sampler_shard")
        return out


def sampler_sizes(n: int, world: int, drop_last: bool) -> tuple[int, int]:
    num, total = c_int64(), c_int64()
    check(_lib.lib().b200ddp_sampler_sizes(int(n), int(world), int(bool(drop_last)), byref(num), byref(total)),
          "sampler_sizes")
    return int(num.value), int(total.value)


def sampler_shard_static(out: torch.Tensor, n, world, rank, seed=0, epoch=0, shuffle=True, drop_last=False,
                         stream=None) -> torch.Tensor:
    """Allocation-free sampler shard (b200ddp_sampler_shard_static): fills the
    caller's int64 CUDA tensor `out` (at least num_samples long, on the current
    device) and returns the filled prefix.  For callers that cannot let the
    library allocate - see include/b200ddp.h.  n is limited to
    _lib.SAMPLER_STATIC_MAX_N; all calls on a device share one scratch array, so
    keep them on one stream."""
    _require_cuda()
    num = sampler_sizes(n, world, drop_last)[0]
    fits = n <= _lib.SAMPLER_STATIC_MAX_N     # beyond the limit the library's own refusal is the error to show
    if fits and (out.dtype != torch.int64 or not out.is_cuda or not out.is_contiguous() or out.numel() < max(num, 0)):
        raise B200DDPError(f"out must be a contiguous int64 CUDA tensor with at least {num} elements")
    if fits and out.is_cuda and out.device.index != torch.cuda.current_device():
        raise B200DDPError(f"out lives on cuda:{out.device.index} but the current device is "
                           f"cuda:{torch.cuda.current_device()}: the entry point works on the current device")
    check(_lib.lib().b200ddp_sampler_shard_static(int(n), int(world), int(rank), int(seed) & 0xFFFFFFFFFFFFFFFF,
                                                  int(epoch) & 0xFFFFFFFFFFFFFFFF, int(bool(shuffle)),
                                                  int(bool(drop_last)),
                                                  c_void_p(out.data_ptr()) if fits and num > 0 else None,
                                                  _stream_handle(stream)), "sampler_shard_static")
    return out[: max(num, 0)]


def sgd_args(lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False, maximize=False,
             first_step=False, grad_divisor=1.0) -> SgdArgs:
    return SgdArgs(float(lr), float(momentum), float(dampening), float(weight_decay), float(grad_divisor),
                   int(bool(nesterov)), int(bool(maximize)), int(bool(first_step)))


def opt_args(rule, lr, beta1=0.9, beta2=0.999, eps=1e-8, grad_divisor=1.0) -> OptArgs:
    """rule: "adam" | "gd" - the lab's own optimizers (codes/task1/pytorch/MyOptimizer.py:18-43)."""
    code = {"adam": _lib.RULE_ADAM, "gd": _lib.RULE_GD}[rule]
    return OptArgs(code, float(lr), float(beta1), float(beta2), float(eps), float(grad_divisor))


# ---- per-process default context ---------------------------------------------------
_default: Context | None = None


def default_context() -> Context:
    """Created on first use for the current device and the default process group."""
    global _default
    if _default is None:
        _require_cuda()
        if dist.is_available() and dist.is_initialized():
            rank, world = dist.get_rank(), dist.get_world_size()
        else:
            rank, world = 0, 1
        if world > _lib.MAX_WORLD:
            raise B200DDPError(f"world size {world} exceeds the single-box limit of {_lib.MAX_WORLD} GPUs")
        ctx = Context(rank, world)
        ctx.connect()
        _default = ctx
    return _default


def reset_default_context():
    global _default
    if _default is not None:
        _default.close()
    _default = None
