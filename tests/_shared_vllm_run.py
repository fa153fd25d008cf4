"""Several CPU tests each need a fresh interpreter with vLLM imported (~12 s a piece).  Their
scripts do not interfere, so they run back to back in ONE subprocess, once per pytest session;
each test then looks for its own marker line.  A script that fails stops the rest: the tests
whose marker is missing report the captured stderr."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_CACHE: dict = {}


def run_all() -> dict:
    if not _CACHE:
        import test_communicator_dispatch as t_dispatch
        import test_model_patch_cpu as t_patch
        import test_norm_provider_cpu as t_norm
        # the dispatch script replaces CudaCommunicator's methods: last
        parts = [t_norm.CODE % ROOT, t_patch.CODE % ROOT, t_dispatch.CODE % ROOT]
        out = subprocess.run([sys.executable, "-c", "\n".join(parts)], capture_output=True, text=True,
                             timeout=900)
        _CACHE.update(rc=out.returncode, out=out.stdout, err=out.stderr)
    return _CACHE
