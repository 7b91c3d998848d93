"""ctypes binding of libpilott_b200.so (the C ABI in include/pilott_b200.h).

There is no CPU fallback: if the library is missing, or an entry point fails, this module
raises.  Build it with ``python -m pilottai_b200.build`` (or ``__graft_entry__.build()``).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PILOTT_B200_LIB") or os.path.join(_HERE, "libpilott_b200.so")   # override: A/B runs of two builds

OK, EINVAL, ECUDA, ENOMEM, EUNSUPPORTED = 0, -1, -2, -3, -4
MAX_FUSED_K = 32
MAX_TAGS = 64
NOALLOC_SCRATCH_BYTES = 16 << 20

c_i64, c_i32, c_u64, c_vp = C.c_int64, C.c_int32, C.c_uint64, C.c_void_p

# name -> (restype, argtypes); every symbol include/pilott_b200.h declares
SIGNATURES = {
    "pilott_last_error": (C.c_char_p, []),
    "pilott_version": (C.c_int, []),
    "pilott_kp": (C.c_int, []),
    "pilott_mem_create": (C.c_int, [C.c_int, c_i64, C.c_int, C.POINTER(c_vp)]),
    "pilott_mem_destroy": (C.c_int, [c_vp]),
    "pilott_mem_clear": (C.c_int, [c_vp, c_vp]),
    "pilott_mem_info": (C.c_int, [c_vp, C.POINTER(c_i64), C.POINTER(c_i64), C.POINTER(c_i64),
                                  C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "pilott_mem_set_option": (C.c_int, [c_vp, C.c_int, c_i64]),
    "pilott_mem_launch_count": (c_i64, [c_vp]),
    "pilott_mem_flagged_total": (C.c_int, [c_vp, C.POINTER(c_i64), C.c_int]),
    "pilott_mem_profile_read": (C.c_int, [c_vp, C.POINTER(C.c_double), C.POINTER(c_i64)]),
    "pilott_mem_append": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, C.POINTER(c_i64), c_vp]),
    "pilott_mem_append_host": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, C.POINTER(c_i64)]),
    "pilott_mem_search": (C.c_int, [c_vp, c_vp, c_i64, C.c_int, c_i32, c_u64, c_i64, c_vp, c_vp, c_vp,
                                    c_vp, c_vp, c_vp]),
    "pilott_mem_search_host": (C.c_int, [c_vp, c_vp, c_i64, C.c_int, c_i32, c_u64, c_i64, c_vp, c_vp,
                                         c_vp, C.POINTER(c_i32)]),
    "pilott_mem_search_host_begin": (C.c_int, [c_vp, C.c_int, c_vp, c_i64, C.c_int, c_i32, c_u64, c_i64]),
    "pilott_mem_search_host_end": (C.c_int, [c_vp, C.c_int, c_vp, c_vp, c_vp, C.POINTER(c_i32)]),
    "pilott_mem_search_large": (C.c_int, [c_vp, c_vp, c_i64, C.c_int, c_i32, c_u64, c_i64, c_vp, c_vp, c_vp,
                                          C.POINTER(c_i32), c_vp]),
    "pilott_mem_search_exact": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, C.c_int, c_i32, c_u64, c_i64,
                                          c_vp, c_vp, c_vp, c_vp]),
    "pilott_mem_search_refine": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, C.c_int, c_i32, c_u64, c_i64,
                                           c_vp, c_vp, c_vp, C.POINTER(c_i32), c_vp]),
    "pilott_mem_search_exact_noalloc": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, C.c_int, c_i32, c_u64, c_i64,
                                                  c_vp, c_vp, c_vp, c_vp]),
    "pilott_mem_compact": (C.c_int, [c_vp, c_i64, C.POINTER(c_i64), c_vp]),
    "pilott_mem_export": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "pilott_xchg_create": (C.c_int, [C.c_int, C.c_int, C.c_int, c_i64, C.POINTER(c_vp), c_vp]),
    "pilott_xchg_connect": (C.c_int, [c_vp, c_vp]),
    "pilott_xchg_publish": (C.c_int, [c_vp, c_vp, c_i64, c_i64, c_vp]),
    "pilott_xchg_publish_copy": (C.c_int, [c_vp, c_vp, c_i64, c_i64, C.c_int, c_vp]),
    "pilott_xchg_wait_buffer": (C.c_int, [c_vp, C.c_int, c_vp]),
    "pilott_xchg_merge": (C.c_int, [c_vp, c_i64, c_i64, C.c_int, c_vp, c_vp, c_vp, c_vp]),
    "pilott_mem_search_publish": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, C.c_int, c_i32, c_u64, c_i64, c_vp, c_vp,
                                            c_vp, c_vp, c_vp, c_vp]),
    "pilott_mem_search_sharded": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, C.c_int, c_i32, c_u64, c_i64, c_vp, c_vp,
                                            c_vp, c_vp, c_vp, c_vp, c_vp]),
    "pilott_xchg_check": (C.c_int, [c_vp]),
    "pilott_xchg_destroy": (C.c_int, [c_vp]),
    "pilott_topk_merge": (C.c_int, [c_vp, c_vp, C.c_int, c_i64, C.c_int, c_vp, c_vp, c_vp, c_vp]),
    "pilott_topk_merge_packed": (C.c_int, [c_vp, C.c_int, c_i64, C.c_int, c_vp, c_vp, c_vp, c_vp]),
    "pilott_route_score": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, C.c_int,
                                     C.c_int, c_vp, C.c_int, C.c_int, c_vp, c_vp, c_vp]),
    "pilott_route_tc_workspace": (c_i64, [c_i64, C.c_int, C.c_int]),
    "pilott_route_score_tc": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, C.c_int,
                                        C.c_int, c_vp, C.c_int, C.c_int, c_vp, c_vp, c_vp, c_i64, c_vp]),
    "pilott_task_priority": (C.c_int, [c_vp, c_vp, c_vp, c_i64, c_vp, c_vp]),
}

_lib = None


class PilottError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"pilott_b200 error {code}: {message}")
        self.code = code


def load():
    """Load the library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA library has not been built "
            "(python -m pilottai_b200.build).  pilottai_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if a declared symbol is not exported
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        msg = load().pilott_last_error()
        raise PilottError(rc, msg.decode() if msg else "unknown error")


def ptr(t) -> int:
    """Device / host address of a torch tensor or numpy array (None -> NULL)."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        return t.data_ptr()
    return t.ctypes.data
