"""Benchmark infrastructure, not product code: lets the STOCK FlashInfer comparator load the
module that was cross-compiled ahead of time into baseline/_fi/aot instead of JIT-compiling it
for minutes on the GPU box.  FlashInfer looks for a package of this name
(flashinfer/jit/env.py:_get_aot_dir); tools/sweep.py and tools/vllm_e2e.py put this directory
on PYTHONPATH when baseline/_fi/aot exists.  Nothing in vllm_distributed_b200 imports it."""
import os

try:
    from flashinfer.version import __version__ as _v
except Exception:  # noqa: BLE001
    _v = "0.0.0"
__version__ = _v

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))


def get_jit_cache_dir() -> str:
    return os.path.join(_ROOT, "baseline", "_fi", "aot")
