"""ctypes front-end to oracle/liboracle.so (see oracle.c for the reference citations).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs; never by vllm_distributed_b200/.
Parity status: pinned against vLLM-generated golden vectors (tests/golden/).
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None

DTYPE_CODE = {torch.bfloat16: 0, torch.float16: 1, torch.float32: 2}


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])
    return _LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.oracle_all_reduce.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int,
                                           ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
        _lib.oracle_fused_add_rms_norm.argtypes = [
            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_float, ctypes.c_int,
            ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        _lib.oracle_all_reduce_add_rms_norm.argtypes = [
            ctypes.POINTER(ctypes.c_void_p), ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
            ctypes.c_float, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
            ctypes.c_void_p, ctypes.c_void_p]
        _lib.oracle_all_gather_lastdim.argtypes = [
            ctypes.POINTER(ctypes.c_void_p), ctypes.c_int, ctypes.c_int, ctypes.c_int,
            ctypes.c_int, ctypes.c_void_p]
        _lib.oracle_pp_slice.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int,
                                         ctypes.c_int, ctypes.c_void_p]
        _lib.oracle_pp_join.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t,
                                        ctypes.c_int, ctypes.c_void_p]
        _lib.oracle_set_num_threads.argtypes = [ctypes.c_int]
    return _lib


def _ptrs(tensors):
    arr = (ctypes.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])
    return arr


def _cpu(t: torch.Tensor) -> torch.Tensor:
    return t.detach().to("cpu").contiguous()


def set_num_threads(n: int) -> None:
    lib().oracle_set_num_threads(int(n))


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def all_reduce(xs: list[torch.Tensor]) -> torch.Tensor:
    xs = [_cpu(x) for x in xs]
    out = torch.empty_like(xs[0])
    rc = lib().oracle_all_reduce(_ptrs(xs), len(xs), xs[0].numel(), DTYPE_CODE[xs[0].dtype],
                                 out.data_ptr())
    assert rc == 0
    return out


def fused_add_rms_norm(x, residual, gamma, eps):
    """-> (norm_out, residual_out); residual may be None (plain rms_norm)."""
    x = _cpu(x)
    residual = None if residual is None else _cpu(residual)
    gamma = None if gamma is None else _cpu(gamma)
    tokens, hidden = x.reshape(-1, x.shape[-1]).shape
    y = torch.empty_like(x)
    r = torch.empty_like(x) if residual is not None else None
    rc = lib().oracle_fused_add_rms_norm(
        x.data_ptr(), None if residual is None else residual.data_ptr(),
        None if gamma is None else gamma.data_ptr(), float(eps), tokens, hidden,
        DTYPE_CODE[x.dtype], y.data_ptr(), None if r is None else r.data_ptr())
    assert rc == 0
    return y, r


def all_reduce_add_rms_norm(xs, residual, gamma, eps):
    """-> (norm_out, residual_out, ar) for the AR -> add -> RMSNorm chain."""
    xs = [_cpu(x) for x in xs]
    residual, gamma = _cpu(residual), _cpu(gamma)
    tokens, hidden = xs[0].shape
    ar, y, r = (torch.empty_like(xs[0]) for _ in range(3))
    rc = lib().oracle_all_reduce_add_rms_norm(
        _ptrs(xs), len(xs), residual.data_ptr(), gamma.data_ptr(), float(eps), tokens, hidden,
        DTYPE_CODE[xs[0].dtype], ar.data_ptr(), y.data_ptr(), r.data_ptr())
    assert rc == 0
    return y, r, ar


def all_gather_lastdim(slabs: list[torch.Tensor]) -> torch.Tensor:
    slabs = [_cpu(s) for s in slabs]
    rows, cols = slabs[0].reshape(-1, slabs[0].shape[-1]).shape
    out = torch.empty(slabs[0].shape[:-1] + (cols * len(slabs),), dtype=slabs[0].dtype)
    rc = lib().oracle_all_gather_lastdim(_ptrs(slabs), len(slabs), rows, cols,
                                         slabs[0].element_size(), out.data_ptr())
    assert rc == 0
    return out


def pp_slice(full: torch.Tensor, tp: int, tp_rank: int) -> torch.Tensor:
    full = _cpu(full)
    out = torch.empty(full.numel() // tp, dtype=full.dtype)
    rc = lib().oracle_pp_slice(full.data_ptr(), full.numel() * full.element_size(), tp, tp_rank,
                               out.data_ptr())
    assert rc == 0
    return out


def pp_join(slices: list[torch.Tensor], shape) -> torch.Tensor:
    slices = [_cpu(s) for s in slices]
    out = torch.empty(shape, dtype=slices[0].dtype)
    rc = lib().oracle_pp_join(_ptrs(slices), out.numel() * out.element_size(), len(slices),
                              out.data_ptr())
    assert rc == 0
    return out


def from_bits(a: np.ndarray) -> torch.Tensor:
    """uint16 bit patterns (golden fixtures) -> bf16 tensor."""
    return torch.from_numpy(a.view(np.int16).copy()).view(torch.bfloat16)


def to_bits(t: torch.Tensor) -> np.ndarray:
    return _cpu(t).view(torch.int16).numpy().view(np.uint16)
