"""ctypes binding of libb200c.so (the C ABI declared in include/b200c.h).

There is no CPU fallback: if the library is missing this module raises, loudly.
Loading the library touches no CUDA state (fork safety for the reference launcher's
local workers, /root/reference/src/launch.py:153-190).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_longlong, c_size_t, c_ubyte, c_void_p

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200C_LIB", os.path.join(HERE, "libb200c.so"))

MAX_RANKS = 8
IPC_HANDLE_BYTES = 64
BF16, F16, F32 = 0, 1, 2
AUTO, ONESHOT, TWOSHOT, NVLS = 0, 1, 2, 3
OK, EINVAL, ECUDA, ETOOBIG, ETIMEOUT, ENOMEM = 0, -1, -2, -3, -4, -5
TUNE_ONESHOT_MAX_BYTES, TUNE_ONESHOT_BLOCKS, TUNE_TWOSHOT_BLOCKS = 1, 2, 3
TUNE_SPIN_TIMEOUT_MS, TUNE_COPY_BLOCKS = 4, 5
TUNE_LINK_BLOCKS, TUNE_PDL, TUNE_ONESHOT_SPLIT, TUNE_DEBUG_EPOCH = 6, 7, 8, 9
TUNE_TRAP_ON_TIMEOUT, TUNE_NVLS_MIN_BYTES, TUNE_NVLS_LINK_BLOCKS, TUNE_NVLS_GROUPS = 10, 11, 12, 13
TUNE_NVLS_COPY_BLOCKS, TUNE_NVLS_TEAMS, TUNE_NVLS_DEBUG = 14, 15, 16
NOALLOC_MAX_BYTES = 64 * 1024

Handle = c_ubyte * IPC_HANDLE_BYTES
PtrArray = POINTER(c_void_p)

# name -> (restype, argtypes); mirrors include/b200c.h one to one
SIGNATURES = {
    "b200c_version": (c_int, []),
    "b200c_last_error": (c_char_p, []),
    "b200c_arena_bytes": (c_int, [c_int, c_size_t, c_size_t, c_size_t, POINTER(c_size_t)]),
    "b200c_alloc_shared": (c_int, [c_size_t, POINTER(c_void_p), POINTER(c_ubyte)]),
    "b200c_open_handle": (c_int, [POINTER(c_ubyte), POINTER(c_void_p)]),
    "b200c_close_handle": (c_int, [c_void_p]),
    "b200c_free_shared": (c_int, [c_void_p]),
    "b200c_init": (c_int, [c_int, c_int, PtrArray, c_size_t, c_size_t, c_size_t, POINTER(c_void_p)]),
    "b200c_destroy": (c_int, [c_void_p]),
    "b200c_nvls_bytes": (c_int, [c_size_t, POINTER(c_size_t)]),
    "b200c_attach_nvls": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t]),
    "b200c_set_tuning": (c_int, [c_void_p, c_int, c_longlong]),
    "b200c_get_tuning": (c_int, [c_void_p, c_int, POINTER(c_longlong)]),
    "b200c_query_error": (c_int, [c_void_p]),
    "b200c_all_reduce": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_int, c_void_p]),
    "b200c_all_reduce_add_rmsnorm": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                             c_void_p, c_float, c_int, c_int, c_int, c_int,
                                             c_void_p]),
    "b200c_fused_add_rms_norm": (c_int, [c_void_p, c_void_p, c_void_p, c_float, c_int, c_int,
                                         c_int, c_void_p]),
    "b200c_rms_norm": (c_int, [c_void_p, c_void_p, c_void_p, c_float, c_int, c_int, c_int,
                               c_void_p]),
    "b200c_all_gather_lastdim": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_size_t, c_void_p]),
    "b200c_send": (c_int, [c_void_p, c_int, c_void_p, c_size_t, c_void_p]),
    "b200c_recv": (c_int, [c_void_p, c_int, c_void_p, c_size_t, c_void_p]),
    "b200c_emu_all_reduce": (c_int, [PtrArray, c_int, PtrArray, PtrArray, c_size_t, c_int, c_int,
                                     c_void_p]),
    "b200c_emu_all_reduce_add_rmsnorm": (c_int, [PtrArray, c_int, PtrArray, PtrArray, PtrArray,
                                                 PtrArray, c_void_p, c_float, c_int, c_int, c_int,
                                                 c_int, c_void_p]),
    "b200c_emu_all_gather_lastdim": (c_int, [PtrArray, c_int, PtrArray, PtrArray, c_int, c_size_t,
                                             c_void_p]),
    "b200c_emu_sendrecv": (c_int, [PtrArray, c_int, c_int, c_int, c_void_p, c_void_p, c_size_t,
                                   c_void_p]),
    "b200c_emu_all_reduce_noalloc": (c_int, [c_int, PtrArray, PtrArray, c_size_t, c_int, c_int,
                                             c_void_p]),
}


class B200CError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libb200c error {code}: {msg}")
        self.code = code


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -m vllm_distributed_b200.build` "
                "(nvcc, sm_100a). There is no CPU or PyTorch fallback for this path.")
        _lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(_lib, name)
            fn.restype = res
            fn.argtypes = args
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise B200CError(rc, lib().b200c_last_error().decode(errors="replace"))


def ptr_array(ptrs) -> "ctypes.Array":
    return (c_void_p * len(ptrs))(*[c_void_p(int(p) if p else 0) for p in ptrs])
