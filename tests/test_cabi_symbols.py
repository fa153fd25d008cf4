"""The C-ABI library loads on a CPU-only box and exports every symbol include/b200c.h declares.
No compute entry point is called here (there is no GPU)."""
import ctypes
import os
import re

from vllm_distributed_b200 import _cabi as C

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200c.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200c_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_bound_and_exported():
    names = declared_symbols()
    assert len(names) >= 20
    lib = C.lib()
    for n in names:
        assert n in C.SIGNATURES, f"{n} declared in b200c.h but not bound in _cabi.py"
        assert hasattr(lib, n), f"{n} not exported by libb200c.so"
    assert sorted(C.SIGNATURES) == names


def test_no_cuda_needed_for_metadata_calls():
    lib = C.lib()
    assert lib.b200c_version() == 1
    out = ctypes.c_size_t(0)
    assert lib.b200c_arena_bytes(8, 1 << 20, 1 << 20, 1 << 20, ctypes.byref(out)) == 0
    # ctrl + 3 regions of 2 halves (max + the 256-byte-aligned chunk-rounding slack) + 2 ag
    # + 8*2 mailboxes
    assert out.value == 64 * 1024 + 6 * ((1 << 20) + 256) + 2 * (1 << 20) + 16 * (1 << 20)
    assert lib.b200c_nvls_bytes(1 << 20, ctypes.byref(out)) == 0
    assert out.value == 32 * 1024 + (1 << 20) + 256  # two flag tables + xin (sums are written in place)
    assert lib.b200c_arena_bytes(9, 16, 16, 16, ctypes.byref(out)) == C.EINVAL
    assert b"world" in lib.b200c_last_error()


def test_header_constants_match_binding():
    text = open(os.path.join(ROOT, "include", "b200c.h")).read()
    consts = dict(re.findall(r"#define\s+B200C_([A-Z0-9_]+)\s+\(?(-?\d+)\)?", text))
    assert int(consts["MAX_RANKS"]) == C.MAX_RANKS
    assert int(consts["IPC_HANDLE_BYTES"]) == C.IPC_HANDLE_BYTES
    # every variant, error code and tuning key the header defines is bound under the same name
    # with the same value (a key added to b200c.h without its binding fails here)
    checked = 0
    for k, v in consts.items():
        if k in ("MAX_RANKS", "IPC_HANDLE_BYTES", "H_", "NOALLOC_MAX_BYTES"):
            continue
        assert hasattr(C, k), f"B200C_{k} is defined in b200c.h but not bound in _cabi.py"
        assert int(v) == getattr(C, k), k
        checked += 1
    assert checked >= 29
    for k in ("NVLS", "TUNE_TRAP_ON_TIMEOUT", "TUNE_NVLS_MIN_BYTES", "TUNE_NVLS_DEBUG", "ETIMEOUT"):
        assert k in consts
