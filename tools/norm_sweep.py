#!/usr/bin/env python3
"""TP=1 leg of the sweep: fused residual-add + RMSNorm alone, HBM-bound at large token counts.
Reports (4*S + 2*H) / t against the measured HBM copy bandwidth, next to vLLM's own
`_C.fused_add_rms_norm` on the same tensors.  One GPU."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from vllm_distributed_b200 import comm as bc  # noqa: E402


def timeit(fn, iters):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        fn(0)
    torch.cuda.synchronize()
    tot = 0.0
    for i in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn(i)
        b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    return tot * 1e3 / iters


def main():
    try:
        peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    have_vllm = False
    try:
        import vllm._custom_ops  # noqa: F401
        have_vllm = hasattr(torch.ops._C, "fused_add_rms_norm")
    except Exception:
        pass
    rows = []
    for H in (4096, 8192):
        gamma = torch.randn(H, device="cuda").bfloat16()
        for T in (64, 256, 1024, 4096, 8192, 16384):
            x = [torch.randn(T, H, device="cuda").bfloat16() for _ in range(2)]
            r = [torch.randn(T, H, device="cuda").bfloat16() for _ in range(2)]
            S = T * H * 2
            us = timeit(lambda i: bc.fused_add_rms_norm_(x[i % 2], r[i % 2], gamma, 1e-5), 10)
            row = {"hidden": H, "tokens": T, "bytes": S, "b200c_us": round(us, 2),
                   "b200c_gbs": round((4 * S + 2 * H) / us / 1e3, 1),
                   "frac_of_measured_hbm": round((4 * S + 2 * H) / us / 1e3 / peak, 3)}
            if have_vllm:
                us2 = timeit(lambda i: torch.ops._C.fused_add_rms_norm(x[i % 2], r[i % 2], gamma, 1e-5), 10)
                row["vllm_C_us"] = round(us2, 2)
                row["vllm_C_gbs"] = round((4 * S + 2 * H) / us2 / 1e3, 1)
            rows.append(row)
            print(json.dumps(row), flush=True)
    out = os.path.join(os.path.dirname(__file__), "..", "gpurun_out", "norm_sweep.json")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    json.dump({"hbm_peak_gbs": peak, "rows": rows}, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()
