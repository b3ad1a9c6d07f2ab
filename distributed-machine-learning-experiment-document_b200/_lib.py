"""ctypes binding of libb200ddp.so (C ABI: include/b200ddp.h).

Marshalling only - no arithmetic happens on this side.  There is no CPU
fallback: if the library is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import (POINTER, Structure, byref, c_char_p, c_double, c_float, c_int, c_int64, c_longlong,
                    c_size_t, c_uint64, c_void_p)

F32, BF16 = 0, 1
DIV_IEEE, DIV_RECIPROCAL = 0, 1
(OPT_DIV_MODE, OPT_PDL, OPT_NVLS_MIN_BYTES, OPT_ONESHOT_MAX_BYTES, OPT_PEER_TIMEOUT_MS, OPT_P2P_BLOCKS,
 OPT_NVLS_DEPTH, OPT_COPY_UNROLL, OPT_TRACE, OPT_HBM_BLOCKS_PER_SM, OPT_NVLS_BLOCKS) = range(1, 12)
OPTIONS = {"div_mode": OPT_DIV_MODE, "pdl": OPT_PDL, "nvls_min_bytes": OPT_NVLS_MIN_BYTES,
           "oneshot_max_bytes": OPT_ONESHOT_MAX_BYTES, "peer_timeout_ms": OPT_PEER_TIMEOUT_MS,
           "p2p_blocks": OPT_P2P_BLOCKS, "nvls_depth": OPT_NVLS_DEPTH, "copy_unroll": OPT_COPY_UNROLL,
           "trace": OPT_TRACE, "hbm_blocks_per_sm": OPT_HBM_BLOCKS_PER_SM,
           "nvls_blocks": OPT_NVLS_BLOCKS}
RULE_ADAM, RULE_GD = 1, 2
MAX_BLOCKS_P2P = 1024
ALGO_AUTO, ALGO_NCCL, ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_NVLS = 0, 1, 2, 3, 4
ALGO_NAMES = {"auto": ALGO_AUTO, "nccl": ALGO_NCCL, "oneshot": ALGO_ONESHOT, "twoshot": ALGO_TWOSHOT,
              "nvls": ALGO_NVLS}
MAX_WORLD = 8
IPC_BYTES = 128
NCCL_UID_BYTES = 128
BUCKET_ALIGN = 32

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200ddp.so")


class B200DDPError(RuntimeError):
    """A call into libb200ddp.so returned an error code."""


class SgdArgs(Structure):
    _fields_ = [("lr", c_double), ("momentum", c_double), ("dampening", c_double),
                ("weight_decay", c_double), ("grad_divisor", c_double),
                ("nesterov", c_int), ("maximize", c_int), ("first_step", c_int)]


class OptArgs(Structure):
    _fields_ = [("rule", c_int), ("lr", c_double), ("beta1", c_double), ("beta2", c_double), ("eps", c_double),
                ("grad_divisor", c_double)]


_PP = POINTER(c_void_p)

_SIGNATURES = {
    "b200ddp_last_error": (c_char_p, []),
    "b200ddp_version": (c_int, []),
    "b200ddp_create": (c_int, [c_int, c_int, c_int, c_size_t, POINTER(c_void_p)]),
    "b200ddp_destroy": (c_int, [c_void_p]),
    "b200ddp_set_option": (c_int, [c_void_p, c_int, c_longlong]),
    "b200ddp_get_option": (c_int, [c_void_p, c_int, POINTER(c_longlong)]),
    "b200ddp_trace_read": (c_int, [c_void_p, POINTER(c_uint64), c_int]),
    "b200ddp_staging": (c_int, [c_void_p, POINTER(c_void_p), POINTER(c_size_t)]),
    "b200ddp_flags": (c_int, [c_void_p, POINTER(c_void_p)]),
    "b200ddp_peer_staging": (c_int, [c_void_p, c_int, POINTER(c_void_p)]),
    "b200ddp_nccl_unique_id": (c_int, [c_void_p]),
    "b200ddp_nccl_init": (c_int, [c_void_p, c_void_p]),
    "b200ddp_ipc_export": (c_int, [c_void_p, c_void_p]),
    "b200ddp_ipc_open_peers": (c_int, [c_void_p, c_void_p]),
    "b200ddp_set_peers_direct": (c_int, [c_void_p, _PP, _PP]),
    "b200ddp_adopt_symmetric": (c_int, [c_void_p, _PP, c_void_p, c_size_t]),
    "b200ddp_free_retired": (c_int, [c_void_p]),
    "b200ddp_close_peers": (c_int, [c_void_p]),
    "b200ddp_plan_create": (c_int, [c_void_p, c_int, POINTER(c_int64), c_int, POINTER(c_void_p)]),
    "b200ddp_plan_destroy": (c_int, [c_void_p]),
    "b200ddp_plan_bucket_elems": (c_int64, [c_void_p]),
    "b200ddp_plan_offsets": (c_int, [c_void_p, POINTER(c_int64)]),
    "b200ddp_pack_scale": (c_int, [c_void_p, c_void_p, _PP, c_void_p, c_float, c_void_p]),
    "b200ddp_allreduce": (c_int, [c_void_p, c_size_t, c_int64, c_int, c_int, c_void_p, POINTER(c_size_t)]),
    "b200ddp_barrier": (c_int, [c_void_p, c_void_p]),
    "b200ddp_allreduce_pick": (c_int, [c_void_p, c_int64, c_int]),
    "b200ddp_unpack": (c_int, [c_void_p, c_void_p, c_void_p, _PP, c_float, c_void_p]),
    "b200ddp_unpack_sgd": (c_int, [c_void_p, c_void_p, c_void_p, _PP, _PP, _PP, POINTER(SgdArgs), c_void_p]),
    "b200ddp_sgd_step": (c_int, [c_void_p, c_void_p, _PP, _PP, _PP, POINTER(SgdArgs), c_void_p]),
    "b200ddp_unpack_opt": (c_int, [c_void_p, c_void_p, c_void_p, _PP, _PP, _PP, _PP, POINTER(OptArgs), c_void_p]),
    "b200ddp_opt_step": (c_int, [c_void_p, c_void_p, _PP, _PP, _PP, _PP, POINTER(OptArgs), c_void_p]),
    "b200ddp_allreduce_sgd": (c_int, [c_void_p, c_void_p, c_size_t, _PP, _PP, _PP, POINTER(SgdArgs), c_void_p]),
    "b200ddp_emu_allreduce": (c_int, [_PP, c_int, c_size_t, c_int64, c_int, c_int, c_void_p, POINTER(c_size_t)]),
    "b200ddp_emu_allreduce_sgd": (c_int, [_PP, _PP, c_int, c_size_t, POINTER(_PP), POINTER(_PP), POINTER(_PP),
                                          POINTER(SgdArgs), c_void_p]),
    "b200ddp_sampler_sizes": (c_int, [c_int64, c_int, c_int, POINTER(c_int64), POINTER(c_int64)]),
    "b200ddp_sampler_shard": (c_int, [c_void_p, c_int64, c_int, c_int, c_uint64, c_uint64, c_int, c_int,
                                      c_void_p, c_void_p]),
    "b200ddp_sampler_shard_static": (c_int, [c_int64, c_int, c_int, c_uint64, c_uint64, c_int, c_int,
                                             c_void_p, c_void_p]),
}
SAMPLER_STATIC_MAX_N = 65536

_lib = None


def lib() -> ctypes.CDLL:
    """The loaded library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200DDPError(
                f"{LIB_PATH} is missing - build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or `make -C <package>/csrc`). There is no CPU fallback.")
        handle = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)   # AttributeError if the ABI and this table ever diverge
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def last_error() -> str:
    msg = lib().b200ddp_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise B200DDPError(f"{what or 'b200ddp call'} failed ({rc}): {last_error()}")


def ptr_array(ptrs) -> ctypes.Array:
    """Host array of device pointers (a ready-made ctypes array passes through)."""
    if isinstance(ptrs, ctypes.Array):
        return ptrs
    return (c_void_p * max(len(ptrs), 1))(*ptrs)


__all__ = ["lib", "check", "last_error", "ptr_array", "SgdArgs", "OptArgs", "B200DDPError", "byref",
           "F32", "BF16", "ALGO_AUTO", "ALGO_NCCL", "ALGO_ONESHOT", "ALGO_TWOSHOT", "ALGO_NAMES"]
