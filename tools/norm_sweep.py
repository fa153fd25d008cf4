#!/usr/bin/env python3
"""TP=1 leg of the sweep: fused residual-add + RMSNorm alone, HBM-bound at large token counts.
Reports (4*S + 2*H) / t against the measured HBM copy bandwidth, next to vLLM's own
`_C.fused_add_rms_norm` on the same tensors.  One GPU.

    python tools/norm_sweep.py                       # the library's own choice of kernel
    python tools/norm_sweep.py --variants rows pipe small2 small4   # + forced kernels (tuning)

Two timings per size: `*_us` = eager single launches after an L2 flush (host launch latency
included); `*_graph_us` = per call inside a CUDA graph of back-to-back calls over rotating
buffers larger than L2 in total (how vLLM runs decode).  GB/s and fractions use the graph numbers.
A forced variant runs in a child process (B200C_NORM_VARIANT is read once per process).
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timeit(torch, fn, iters):
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


def time_graph(torch, fn, calls):
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        fn(0)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(st), torch.cuda.graph(g, stream=st):
        for i in range(calls):
            fn(i)
    torch.cuda.synchronize()
    best = None
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(st):
            a.record(st)
            g.replay()
            b.record(st)
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        best = ms if best is None else min(best, ms)
    return best * 1e3 / calls


def measure(args):
    import torch
    from vllm_distributed_b200 import comm as bc
    have_vllm = False
    if args.with_vllm:
        try:
            import vllm._custom_ops  # noqa: F401
            have_vllm = hasattr(torch.ops._C, "fused_add_rms_norm")
        except Exception:  # noqa: BLE001
            pass
    rows = []
    for H in args.hidden:
        gamma = torch.randn(H, device="cuda").bfloat16()
        for T in args.tokens:
            S = T * H * 2
            nbuf = max(2, min(16, (512 << 20) // (4 * S)))
            x = [torch.randn(T, H, device="cuda").bfloat16() for _ in range(nbuf)]
            r = [torch.randn(T, H, device="cuda").bfloat16() for _ in range(nbuf)]
            calls = max(nbuf, 32 if S <= (8 << 20) else nbuf)
            ours = lambda i: bc.fused_add_rms_norm_(x[i % nbuf], r[i % nbuf], gamma, 1e-5)  # noqa: E731
            row = {"hidden": H, "tokens": T, "bytes": S, "us": round(timeit(torch, ours, 10), 2),
                   "graph_us": round(time_graph(torch, ours, calls), 2)}
            if have_vllm:
                theirs = lambda i: torch.ops._C.fused_add_rms_norm(x[i % nbuf], r[i % nbuf], gamma, 1e-5)  # noqa: E731
                row["vllm_C_us"] = round(timeit(torch, theirs, 10), 2)
                row["vllm_C_graph_us"] = round(time_graph(torch, theirs, calls), 2)
            rows.append(row)
            del x, r
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hidden", type=int, nargs="+", default=[4096, 8192])
    ap.add_argument("--tokens", type=int, nargs="+", default=[64, 256, 1024, 2048, 4096, 8192, 16384])
    ap.add_argument("--variants", nargs="*", default=[])
    ap.add_argument("--child", default="")
    ap.add_argument("--with-vllm", type=int, default=1)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "norm_sweep.json"))
    args = ap.parse_args()
    if args.child:
        print("ROWS " + json.dumps(measure(args)), flush=True)
        return
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:  # noqa: BLE001
        peak = 6650.0
    rows = measure(args)
    for row in rows:
        alg = 4 * row["bytes"] + 2 * row["hidden"]
        row["b200c_us"], row["b200c_graph_us"] = row.pop("us"), row.pop("graph_us")
        row["b200c_gbs"] = round(alg / row["b200c_graph_us"] / 1e3, 1)
        row["frac_of_measured_hbm"] = round(alg / row["b200c_graph_us"] / 1e3 / peak, 3)
        if "vllm_C_graph_us" in row:
            row["vllm_C_gbs"] = round(alg / row["vllm_C_graph_us"] / 1e3, 1)
            row["b200c_over_vllm_C"] = round(row["b200c_graph_us"] / row["vllm_C_graph_us"], 3)
    for v in args.variants:
        env = dict(os.environ, B200C_NORM_VARIANT=v)
        cmd = [sys.executable, os.path.abspath(__file__), "--child", v, "--with-vllm", "0",
               "--hidden", *map(str, args.hidden), "--tokens", *map(str, args.tokens)]
        out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
        line = [ln for ln in out.stdout.splitlines() if ln.startswith("ROWS ")]
        if not line:
            print(f"variant {v} failed: {out.stderr[-500:]}", file=sys.stderr)
            continue
        for row, extra in zip(rows, json.loads(line[0][5:])):
            row[f"{v}_graph_us"] = extra["graph_us"]
    for row in rows:
        print(json.dumps(row), flush=True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump({"hbm_peak_gbs": peak,
               "note": "*_us: eager single launches after an L2 flush (host launch latency included); "
                       "*_graph_us: per call inside a CUDA graph of back-to-back calls over rotating "
                       "buffers larger than L2 in total; GB/s and fractions use the graph numbers; "
                       "<variant>_graph_us: the same with B200C_NORM_VARIANT forced (tuning)",
               "rows": rows}, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
