#!/usr/bin/env python3
"""Smallest end-to-end exercise of every kernel family, meant to be run under
`compute-sanitizer --tool memcheck` (one tool per call, see B200_PROFILING.md)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from vllm_distributed_b200 import _cabi as C  # noqa: E402
from vllm_distributed_b200 import comm  # noqa: E402
from vllm_distributed_b200.emu import EmuGroup  # noqa: E402

g = EmuGroup(4, "cuda:0", max_bytes=1 << 20, ag_bytes=1 << 20, p2p_bytes=64 << 10)
g.set_tuning(C.TUNE_SPIN_TIMEOUT_MS, 60000)
for t, h in ((3, 896), (2, 8192), (9, 1000)):
    xs = [torch.randn(t, h, device="cuda").bfloat16() for _ in range(4)]
    res = [torch.randn(t, h, device="cuda").bfloat16() for _ in range(4)]
    gamma = torch.randn(h, device="cuda").bfloat16()
    for variant in (1, 2):
        g.all_reduce(xs, variant=variant)
        g.all_reduce_add_rmsnorm(xs, res, gamma, 1e-5, variant=variant)
g.all_gather_lastdim([torch.randn(5, 264, device="cuda").bfloat16() for _ in range(4)])
g.all_gather_lastdim([torch.randn(5, 2, device="cuda") for _ in range(4)])
for i in range(4):
    g.sendrecv(0, 1, torch.randn(4096, device="cuda").bfloat16())
x, r = torch.randn(7, 4096, device="cuda").bfloat16(), torch.randn(7, 4096, device="cuda").bfloat16()
comm.fused_add_rms_norm_(x, r, torch.randn(4096, device="cuda").bfloat16(), 1e-5)
torch.cuda.synchronize()
g.check_health()
print("sanitize smoke ok")
