#!/usr/bin/env python3
"""Time the emulated N-rank kernels on ONE GPU (all 'link' traffic is local): separates what the
kernels cost by themselves from what the fabric costs.  Also the ncu target for the collective
kernels (ncu must never wrap a multi-rank command)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from vllm_distributed_b200 import _cabi as C  # noqa: E402
from vllm_distributed_b200.emu import EmuGroup  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=2)
    ap.add_argument("--tokens", type=int, nargs="+", default=[16, 64, 256, 1024, 4096])
    ap.add_argument("--hidden", type=int, default=8192)
    ap.add_argument("--variant", default="twoshot")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--link", type=int, default=0)
    args = ap.parse_args()
    n = args.world
    mx = max(args.tokens) * args.hidden * 2
    g = EmuGroup(n, "cuda:0", max_bytes=mx, ag_bytes=1 << 20, p2p_bytes=1 << 20)
    g.set_tuning(C.TUNE_SPIN_TIMEOUT_MS, 5000)
    if args.link:
        g.set_tuning(C.TUNE_LINK_BLOCKS, args.link)
    variant = {"oneshot": C.ONESHOT, "twoshot": C.TWOSHOT, "auto": C.AUTO}[args.variant]
    gamma = torch.randn(args.hidden, device="cuda").bfloat16()
    for T in args.tokens:
        if variant == C.ONESHOT and T * args.hidden * 2 * n > 2 * mx:
            continue
        xs = [torch.randn(T, args.hidden, device="cuda").bfloat16() for _ in range(n)]
        res = [torch.randn(T, args.hidden, device="cuda").bfloat16() for _ in range(n)]
        outs = [torch.empty_like(x) for x in xs]
        for fused in (True, False):
            def call():
                if fused:
                    g.all_reduce_add_rmsnorm(xs, res, gamma, 1e-5, variant=variant, inplace=False)
                else:
                    g.all_reduce(xs, variant=variant, outs=outs)
            for _ in range(3):
                call()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(args.iters):
                call()
            b.record()
            torch.cuda.synchronize()
            us = a.elapsed_time(b) * 1e3 / args.iters
            S = T * args.hidden * 2
            print(f"world={n} T={T} S={S >> 10} KiB {args.variant} {'fused' if fused else 'plain'}: "
                  f"{us:8.1f} us  ({n} ranks on one GPU; per-rank-equivalent HBM ~{9 * S * n / us / 1e6:.0f} GB/s)")
        g.check_health()


if __name__ == "__main__":
    main()
