#!/usr/bin/env python3
"""Bench infrastructure: cross-compile FlashInfer's trtllm_comm module (the reference's fused
all-reduce kernels) for sm_100a here, where there is no GPU, so that the GPU box
does not spend 5 minutes of its lease JIT-compiling the comparator.  Output:
baseline/_fi/aot/trtllm_comm/trtllm_comm.so (git-ignored; travels with gpurun), picked up through
tools/fi_jit_cache_shim by tools/sweep.py and tools/vllm_e2e.py.  Takes about 5 minutes."""
import os
import shutil
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ["FLASHINFER_CUDA_ARCH_LIST"] = "10.0a"
os.environ.setdefault("FLASHINFER_WORKSPACE_BASE", "/tmp/fi_ws")

from flashinfer.jit import env  # noqa: E402
from flashinfer.jit.comm import gen_trtllm_comm_module  # noqa: E402

t = time.time()
spec = gen_trtllm_comm_module()
spec.build(verbose=False)
src = os.path.join(str(env.FLASHINFER_JIT_DIR), "trtllm_comm", "trtllm_comm.so")
dst = os.path.join(ROOT, "baseline", "_fi", "aot", "trtllm_comm")
os.makedirs(dst, exist_ok=True)
shutil.copy2(src, dst)
print(f"built in {time.time() - t:.0f} s -> {dst}/trtllm_comm.so")
