"""ctypes loader for the in-tree CUDA library.

The library is built by ``grace_b200/csrc/build.sh`` (or ``__graft_entry__.build()``) into
``grace_b200/lib/libgrace_b200.so``.  On a machine with a GPU a missing library is an
error, never a silent fallback: ``lib()`` raises and the fused engine refuses to start.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# GRACE_B200_LIB picks another build of the same library, e.g. the -DGB_DEBUG one (bounds asserts in every
# emit / merge / decode store) that ``GB_DEBUG=1 bash grace_b200/csrc/build.sh`` writes to lib/debug/
LIB_PATH = os.environ.get("GRACE_B200_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "libgrace_b200.so")
BUILD_SH = os.path.join(os.path.dirname(_HERE), "csrc", "build.sh")

CODEC = dict(none=0, fp16=1, topk=2, randomk=3, threshold=4, dgc=5, qsgd=6, terngrad=7, natural=8, u8bit=9,
             signsgd=10, signum=11, efsignsgd=12, onebit=13, sketch=14, adaq=15, inceptionn=16)
MEMORY = dict(none=0, residual=1, dgc=2)


class NativeError(RuntimeError):
    pass


class CodecParams(C.Structure):
    """Mirror of ``gb_codec_params``."""
    _fields_ = [("ratio", C.c_double), ("codec", C.c_int32), ("memory", C.c_int32), ("beta", C.c_float),
                ("gamma", C.c_float), ("threshold", C.c_float), ("quantum_num", C.c_int32), ("momentum", C.c_float),
                ("lr", C.c_float), ("seed", C.c_uint32), ("average", C.c_int32), ("capacity", C.c_float),
                ("one_minus_momentum", C.c_float), ("clip", C.c_int32)]


_lib = None


def build(verbose=False):
    out = subprocess.run(["bash", BUILD_SH], capture_output=True, text=True)
    if out.returncode != 0:
        raise NativeError("building libgrace_b200.so failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stdout)
    return LIB_PATH


def available():
    return os.path.exists(LIB_PATH)


def _declare(L):
    vp, i64, u32, i32, sz = C.c_void_p, C.c_int64, C.c_uint32, C.c_int, C.c_size_t
    P = C.POINTER
    sigs = {
        "gb_last_error": (C.c_char_p, []),
        "gb_version": (i32, []),
        "gb_set_device": (i32, [i32]),
        "gb_comm_create": (i32, [i32, i32, sz, P(vp)]),
        "gb_comm_ipc_handle": (i32, [vp, vp]),
        "gb_comm_open_peers": (i32, [vp, vp]),
        "gb_comm_create_local": (i32, [i32, sz, P(vp)]),
        "gb_comm_destroy": (i32, [vp]),
        "gb_comm_status": (i32, [vp, P(u32)]),
        "gb_comm_clear_status": (i32, [vp]),
        "gb_comm_check_guards": (i32, [vp, P(i32)]),
        "gb_comm_guard_poke": (i32, [vp]),
        "gb_comm_set_timeout_ms": (i32, [vp, C.c_uint64]),
        "gb_comm_rank": (i32, [vp]),
        "gb_comm_world": (i32, [vp]),
        "gb_comm_arena_free": (sz, [vp]),
        "gb_comm_barrier": (i32, [vp, vp]),
        "gb_comm_arena_need": (sz, [vp]),
        "gb_comm_regions": (i32, [vp]),
        "gb_comm_region_add": (i32, [vp, sz, P(i32)]),
        "gb_comm_region_adopt": (i32, [vp, P(C.c_uint64), C.c_uint64, sz, P(i32)]),
        "gb_comm_has_multicast": (i32, [vp]),
        "gb_comm_region_ipc_handle": (i32, [vp, i32, vp]),
        "gb_comm_region_open_peers": (i32, [vp, i32, vp]),
        "gb_handle_record": (i32, [vp, P(i64)]),
        "gb_handle_poll": (i32, [i64]),
        "gb_handle_wait": (i32, [i64]),
        "gb_handle_stream_wait": (i32, [i64, vp]),
        "gb_handle_release": (i32, [i64]),
        "gb_ops_create": (i32, [vp, sz, P(vp)]),
        "gb_ops_destroy": (i32, [vp]),
        "gb_ops_capacity": (sz, [vp]),
        "gb_ops_launches": (i64, [vp]),
        "gb_ops_send": (i32, [vp, vp, sz, vp]),
        "gb_ops_recv": (i32, [vp, sz, i32, i32, sz, i32, C.c_double, vp, vp]),
        "gb_plan_create": (i32, [vp, i32, P(i64), P(u32), P(CodecParams), P(vp)]),
        "gb_plan_destroy": (i32, [vp]),
        "gb_plan_padded_numel": (i64, [vp]),
        "gb_plan_offsets": (i32, [vp, P(i64)]),
        "gb_plan_wire_bytes": (i64, [vp]),
        "gb_plan_send": (i32, [vp, vp, vp]),
        "gb_plan_presend": (i32, [vp, vp, vp]),
        "gb_plan_recv": (i32, [vp, vp, vp]),
        "gb_plan_mid": (i32, [vp, vp]),
        "gb_plan_wait": (i32, [vp, vp]),
        "gb_psgd_wait": (i32, [vp, i32, vp]),
        "gb_plan_state": (i32, [vp, i32, P(vp), P(sz)]),
        "gb_plan_state_read": (i32, [vp, i32, vp, vp]),
        "gb_plan_state_write": (i32, [vp, i32, vp, vp]),
        "gb_plan_get_step": (i32, [vp, P(u32)]),
        "gb_plan_set_step": (i32, [vp, u32]),
        "gb_plan_launches": (i32, [vp, P(i64)]),
        "gb_nvls_create": (i32, [vp, i64, P(vp)]),
        "gb_nvls_destroy": (i32, [vp]),
        "gb_nvls_signal": (i32, [vp, vp]),
        "gb_nvls_reduce": (i32, [vp, vp, i32, vp]),
        "gb_nvls_launches": (i32, [vp, P(i64)]),
        "gb_psgd_create": (i32, [vp, i32, P(i64), P(i64), P(u32), i32, u32, i32, i32, P(vp)]),
        "gb_psgd_create2": (i32, [vp, i32, P(i64), P(i64), P(u32), i32, u32, i32, i32, i32, P(vp)]),
        "gb_psgd_qprev": (i32, [vp, P(vp), P(sz)]),
        "gb_psgd_qprev_copy": (i32, [vp, vp, i32, vp]),
        "gb_psgd_destroy": (i32, [vp]),
        "gb_psgd_padded_numel": (i64, [vp]),
        "gb_psgd_offsets": (i32, [vp, P(i64)]),
        "gb_psgd_send": (i32, [vp, vp, vp]),
        "gb_psgd_mid": (i32, [vp, vp]),
        "gb_psgd_recv": (i32, [vp, vp, vp]),
        "gb_psgd_launches": (i32, [vp, P(i64)]),
        "gb_psgd_tc_matrices": (i32, [vp]),
        "gb_psgd_get_step": (i32, [vp, P(u32)]),
        "gb_psgd_set_step": (i32, [vp, u32]),
        "gb_psgd_residual": (i32, [vp, P(vp), P(sz)]),
        "gb_psgd_wire_bytes": (i64, [vp]),
        "gb_psgd_state_read": (i32, [vp, vp, vp]),
        "gb_psgd_state_write": (i32, [vp, vp, vp]),
        "gb_psgd_residual": (i32, [vp, P(vp), P(sz)]),
        "gb_psgd_peek": (i32, [vp, i32, vp, i64, vp]),
        "gb_topk_scratch_bytes": (sz, [i64]),
        "gb_topk_compress": (i32, [vp, i64, i64, vp, vp, vp, sz, vp]),
        "gb_topk_static_max_n": (i64, []),
        "gb_topk_compress_static": (i32, [vp, i64, i64, vp, vp, vp]),
        "gb_sparse_decompress": (i32, [P(vp), P(vp), P(i64), i32, i64, C.c_float, vp, vp]),
    }
    for name, (res, args) in sigs.items():
        if hasattr(L, name):
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
    return L


def lib():
    """Load (once) and return the native library; raises ``NativeError`` if it is absent."""
    global _lib
    if _lib is None:
        if not available():
            raise NativeError("%s not found -- run `bash %s` (or __graft_entry__.build())" % (LIB_PATH, BUILD_SH))
        _lib = _declare(C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL))
    return _lib


def check(rc):
    if rc != 0:
        raise NativeError(lib().gb_last_error().decode())
