#!/usr/bin/env python3
"""bench.py -- the TP hot path of one decode step, measured on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path over one decode batch of synthetic activations:
  N == 1 : Llama-3-8B-shaped, TP=1 -- 2L+1 = 65 fused residual-add + RMSNorm calls on
           [B, 4096] bf16 (the all-reduce is the identity at TP=1, parallel_state.py:530).
  N >= 2 : Llama-3-8B (N=2) / Llama-3-70B (N=4, 8) shaped, TP=N -- 2L+1 fused
           all-reduce + residual-add + RMSNorm calls on [B, H] plus one lm_head logits
           all-gather [B, V/N] -> [B, V].
GEMMs and attention are NOT part of the step: this is the communication / normalisation
hot path the north star names, so "tokens/s" is B / (hot-path time of one decode step).

Rank 0 prints ONE JSON line (contract in the task statement).  `value` is measured with
inputs resident in HBM and the step replayed as a CUDA graph (as vLLM runs decode);
`e2e` goes through the public Python API eagerly with host buffers: every step copies the
step's inputs from pinned host memory and reads the result back.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "decode tokens/s (TP hot path: per-layer all-reduce+residual+RMSNorm, lm_head gather)"
NVLINK_PEAK_GBS = 770.0  # measured peer-copy bandwidth per direction on this pool (B200_PROFILING.md)
HBM_FALLBACK_GBS = 6650.0


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def model_shape(n_gpus: int) -> dict:
    if n_gpus <= 2:
        return dict(model="llama3-8b-shaped", hidden=4096, layers=32, vocab=128256, eps=1e-5)
    return dict(model="llama3-70b-shaped", hidden=8192, layers=80, vocab=128256, eps=1e-5)


def measured_peaks() -> tuple[float, str]:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy)"
    except Exception:
        return HBM_FALLBACK_GBS, "fallback 6.65 TB/s (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx = float(f[2])
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


# =================================== our arm =========================================================

def run_ours(args) -> dict:
    import torch
    import torch.distributed as dist
    from vllm_distributed_b200 import _cabi as C
    from vllm_distributed_b200 import comm as bc

    n = args.gpus
    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if world != n:
        raise SystemExit(f"--gpus {n} but WORLD_SIZE={world}: launch with torch.distributed.run "
                         f"--nproc-per-node {n}")
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback exists)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    C.lib()  # fail loudly if the extension is missing

    shp = model_shape(n)
    H, L, V, eps = shp["hidden"], shp["layers"], shp["vocab"], shp["eps"]
    B = args.batch
    calls = 2 * L + 1
    S = B * H * 2

    cpu_group = None
    if n > 1:
        dist.init_process_group(backend="nccl", device_id=dev)
        cpu_group = dist.new_group(backend="gloo")
        comm = bc.B200Comm.from_process_group(cpu_group, dev, max_bytes=max(S, 1 << 20),
                                              ag_bytes=max(B * V * 2, 1 << 20), p2p_bytes=1 << 20)
    else:
        comm = bc.B200Comm(0, 1, None, 0, 0, 0, dev)

    gen = torch.Generator(device="cpu").manual_seed(1234 + rank)
    # distinct per-layer partial sums (stand-ins for the o_proj / down_proj GEMM outputs)
    x_host = torch.randn(calls, B, H, generator=gen).bfloat16().pin_memory()
    g7 = torch.Generator(device="cpu").manual_seed(7)
    res_host = torch.randn(B, H, generator=g7).bfloat16().pin_memory()
    g8 = torch.Generator(device="cpu").manual_seed(8)
    gammas = torch.randn(calls, H, generator=g8).bfloat16().to(dev)
    logits_host = torch.randn(B, V // n, generator=gen).bfloat16().pin_memory()

    x = x_host.to(dev)
    residual = res_host.to(dev)
    h = torch.empty(B, H, dtype=torch.bfloat16, device=dev)
    logits_local = logits_host.to(dev)
    logits_full = torch.empty(B, V, dtype=torch.bfloat16, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    variant = {"auto": C.AUTO, "oneshot": C.ONESHOT, "twoshot": C.TWOSHOT}[args.variant]

    def step():
        if n == 1:
            for i in range(calls):
                # residual += x[i]; h = rmsnorm(residual) * gamma[i]   (out of place on x)
                comm.all_reduce_add_rmsnorm_into(x[i], residual, gammas[i], eps, h)
        else:
            for i in range(calls):
                comm.all_reduce_add_rmsnorm_into(x[i], residual, gammas[i], eps, h, variant)
            comm.all_gather_lastdim(logits_local, out=logits_full)

    launches_per_step = calls + (1 if n > 1 else 0)

    def sync_all():
        torch.cuda.synchronize()
        if n > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident measurement: the step as a CUDA graph --------------------------------
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        for _ in range(2):
            step()
    sync_all()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(stream):
        with torch.cuda.graph(graph, stream=stream):
            step()
    sync_all()

    def timed_steps(fn, k, w):
        with torch.cuda.stream(stream):
            for _ in range(w):
                flush.zero_()
                fn()
        sync_all()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
               for _ in range(k)]
        sampler = ClockSampler(local_rank)
        if rank == 0:
            sampler.start()
        with torch.cuda.stream(stream):
            for a, b in evs:
                flush.zero_()  # evict the previous step's lines from L2 (untimed)
                a.record(stream)
                fn()
                b.record(stream)
        sync_all()
        clocks = sampler.stop() if rank == 0 else None
        total_ms = sum(a.elapsed_time(b) for a, b in evs)
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        if n > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / k, clocks

    ms_per_step, clocks = timed_steps(graph.replay, args.steps, max(3, args.warmup))

    # ---- end to end: public API, eager, host buffers every step ---------------------------------
    out_host = torch.empty(B, H, dtype=torch.bfloat16).pin_memory()
    logits_out_host = torch.empty(B, V, dtype=torch.bfloat16).pin_memory() if n > 1 else None

    def e2e_step():
        x.copy_(x_host, non_blocking=True)
        residual.copy_(res_host, non_blocking=True)
        if n > 1:
            logits_local.copy_(logits_host, non_blocking=True)
        step()
        out_host.copy_(h, non_blocking=True)
        if n > 1:
            logits_out_host.copy_(logits_full, non_blocking=True)

    e2e_steps = max(3, min(args.steps, 20))
    e2e_ms, _ = timed_steps(e2e_step, e2e_steps, 3)
    h2d = x_host.numel() * 2 + res_host.numel() * 2 + (logits_host.numel() * 2 if n > 1 else 0)
    d2h = out_host.numel() * 2 + (logits_out_host.numel() * 2 if n > 1 else 0)

    comm.check_health()

    # ---- roofline of the dominant kernel -----------------------------------------------------------
    hbm_peak, hbm_src = measured_peaks()
    if n == 1:
        alg_bytes = 4 * S + 2 * H  # read x, read residual, write residual, write norm, read gamma
        achieved = alg_bytes * calls / (ms_per_step * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": "norm_kernel<bf16,1>", "achieved": round(achieved, 2),
                "peak": hbm_peak, "unit": "GB/s", "frac": round(achieved / hbm_peak, 4),
                # dram__bytes_read.sum + dram__bytes_write.sum per launch from the ncu --set full
                # capture of this kernel at this shape (profiles/r1_n1_norm_kernel_full_raw.csv):
                # 1.07 MB read (x + residual), 0 written (the outputs stay in the 126 MB L2)
                "traffic": 1066496 if (B, H) == (64, 4096) else None, "peak_source": hbm_src,
                "algorithmic_bytes_per_launch": alg_bytes}
    else:
        bus_bytes = 2 * (n - 1) / n * S  # all-reduce bus bytes
        achieved = bus_bytes * calls / (ms_per_step * 1e-3) / 1e9
        roof = {"bound": "nvlink", "kernel": "ar_norm_kernel<bf16>", "achieved": round(achieved, 2),
                "peak": NVLINK_PEAK_GBS, "unit": "GB/s", "frac": round(achieved / NVLINK_PEAK_GBS, 4),
                "traffic": None,
                "peak_source": "measured peer copy 770 GB/s per direction (B200_PROFILING.md); "
                               "nominal 900",
                "algorithmic_bytes_per_launch": bus_bytes,
                "note": "achieved counts the step's all-reduce bus bytes over the whole step time "
                        "(the lm_head gather is inside the step)"}

    result = {
        "metric": METRIC,
        "value": round(B / (ms_per_step * 1e-3), 1),
        "unit": "tokens/s",
        "n_gpus": n,
        "steps": args.steps,
        "warmup": max(3, args.warmup),
        "ms_per_step": round(ms_per_step, 5),
        "higher_is_better": True,
        "scaling": "weak",
        "vs_baseline": None,
        "dtype": "bf16",
        "data": "synthetic",
        "config": {
            "workload": f"{shp['model']} TP={n} decode hot path, batch {B}: {calls} x "
                        + ("fused residual-add+RMSNorm" if n == 1 else
                           "fused all-reduce+residual-add+RMSNorm + 1 lm_head all-gather")
                        + f" on [{B},{H}] bf16",
            "hidden": H, "layers": L, "vocab": V, "batch": B, "tp": n,
            "calls_per_step": launches_per_step,
            "l2": "256 MiB buffer rewritten between timed steps (untimed) to flush the 126 MB L2",
            "execution": "step captured once as a CUDA graph, replayed per timed step",
            "variant": args.variant,
        },
        "clocks": clocks,
        "e2e": {"value": round(B / (e2e_ms * 1e-3), 1), "unit": "tokens/s",
                "ms_per_step": round(e2e_ms, 5), "steps": e2e_steps,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "path": "B200Comm python API, eager launches, pinned host buffers"},
        "gpu_launches": launches_per_step * args.steps,
        "roofline": roof,
        "us_per_call": round(ms_per_step * 1e3 / launches_per_step, 3),
    }
    if n == 1 and rank == 0 and not args.no_cpu_baseline:
        result["cpu_baseline"] = cpu_baseline_port(shp, B, calls)
    if n > 1:
        dist.barrier()
        comm.close()
        dist.destroy_process_group()
    return result if rank == 0 else None


# ================================ CPU baselines ==========================================================

def cpu_baseline_port(shp: dict, B: int, calls: int, budget_s: float = 10.0) -> dict:
    """The oracle (a port of the reference's algorithm) timed on the host cores: the same
    step, as many repetitions as fit in ~budget_s."""
    import torch
    from oracle import oracle
    H, eps = shp["hidden"], shp["eps"]
    cores = os.cpu_count() or 1
    oracle.set_num_threads(cores)
    g = torch.Generator().manual_seed(1)
    xs = [torch.randn(B, H, generator=g).bfloat16() for _ in range(4)]
    res = torch.randn(B, H, generator=g).bfloat16()
    gamma = torch.randn(H, generator=g).bfloat16()
    t0 = time.perf_counter()
    steps = 0
    while True:
        for i in range(calls):
            _, res = oracle.fused_add_rms_norm(xs[i % 4], res, gamma, eps)
        steps += 1
        if time.perf_counter() - t0 > budget_s or steps >= 200:
            break
    dt = (time.perf_counter() - t0) / steps
    return {"value": round(B / dt, 1), "unit": "tokens/s", "cores": oracle.num_threads(),
            "kind": "port", "ms_per_step": round(dt * 1e3, 3),
            "sample": f"{steps} full steps ({calls} fused residual-add+RMSNorm calls on [{B},{H}] "
                      f"bf16 each) of oracle/oracle.c with OpenMP"}


def _ref_worker(rank, world, port, shp, B, calls, steps, warmup, use_vllm, q):
    """One rank of the reference's CPU path: vLLM's own communicator API over gloo."""
    import torch
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), CUDA_VISIBLE_DEVICES="")
    H, V, eps = shp["hidden"], shp["vocab"], shp["eps"]
    threads = max(1, (os.cpu_count() or 1) // world)
    torch.set_num_threads(threads)
    kind = "reference"
    try:
        if not use_vllm:
            raise RuntimeError("vllm path disabled")
        from vllm.config import DeviceConfig, VllmConfig, set_current_vllm_config
        from vllm.distributed import parallel_state as ps
        from vllm.distributed.communication_op import (tensor_model_parallel_all_gather,
                                                       tensor_model_parallel_all_reduce)
        import vllm.ir.ops as irops
        ctx = set_current_vllm_config(VllmConfig(device_config=DeviceConfig(device="cpu")))
        ctx.__enter__()
        ps.init_distributed_environment(world_size=world, rank=rank, local_rank=rank,
                                        distributed_init_method=f"tcp://127.0.0.1:{port}",
                                        backend="gloo")
        ps.initialize_model_parallel(world, 1)
        ar = tensor_model_parallel_all_reduce
        ag = tensor_model_parallel_all_gather
        norm = irops.fused_add_rms_norm
    except Exception as e:  # vLLM's CPU path unusable on this box: time the port instead
        kind = f"port ({type(e).__name__})"
        import torch.distributed as dist
        if not dist.is_initialized():
            dist.init_process_group("gloo", rank=rank, world_size=world,
                                    init_method=f"tcp://127.0.0.1:{port}")

        def ar(t):
            if world > 1:
                dist.all_reduce(t)
            return t

        def ag(t, dim=-1):
            if world == 1:
                return t
            outs = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(outs, t)
            return torch.cat(outs, dim=-1)

        def norm(x, r, w, e):
            x32 = x.float() + r.float()
            r2 = x32.to(x.dtype)
            y = (x32 * torch.rsqrt(x32.pow(2).mean(-1, keepdim=True) + e)).to(w.dtype) * w
            return y, r2
    import torch.distributed as dist
    g = torch.Generator().manual_seed(1234 + rank)
    xs = [torch.randn(B, H, generator=g).bfloat16() for _ in range(4)]
    res = torch.randn(B, H, generator=torch.Generator().manual_seed(7)).bfloat16()
    gamma = torch.randn(H, generator=torch.Generator().manual_seed(8)).bfloat16()
    logits = torch.randn(B, V // world, generator=g).bfloat16()

    def step(res):
        for i in range(calls):
            s = ar(xs[i % 4].clone()) if world > 1 else xs[i % 4]
            _, res = norm(s, res, gamma, eps)
        if world > 1:
            ag(logits, -1)
        return res

    for _ in range(warmup):
        res = step(res)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        res = step(res)
    if world > 1:
        dist.barrier()
    dt = (time.perf_counter() - t0) / steps
    q.put((rank, dt, kind, threads))


def run_reference(args) -> dict | None:
    """--impl reference: the reference's own CPU implementation of the path (vLLM's
    communicator API over gloo + its native fused_add_rms_norm) on the box's host cores.
    Under torchrun only rank 0 works; it spawns its own N gloo workers."""
    if int(os.environ.get("RANK", 0)) != 0:
        return None
    import torch.multiprocessing as mp
    n = args.gpus
    shp = model_shape(n)
    B, calls = args.batch, 2 * shp["layers"] + 1
    steps = max(1, min(args.steps, 10))
    warmup = max(1, min(args.warmup, 2))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29400 + (os.getpid() % 500)
    procs = [ctx.Process(target=_ref_worker,
                         args=(r, n, port, shp, B, calls, steps, warmup, not args.no_vllm, q))
             for r in range(n)]
    for p in procs:
        p.start()
    outs = [q.get(timeout=1500) for _ in range(n)]
    for p in procs:
        p.join(timeout=60)
    dt = max(o[1] for o in outs)
    kind = outs[0][2]
    threads = outs[0][3]
    val = round(B / dt, 1)
    sample = (f"{steps} full steps of the same workload, {n} gloo process(es) x {threads} threads")
    return {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "tokens/s", "n_gpus": n,
        "steps": steps, "warmup": warmup, "ms_per_step": round(dt * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": {"workload": f"{shp['model']} TP={n} decode hot path, batch {B} on host cores "
                               f"(vLLM CPU/gloo path)", "hidden": shp["hidden"],
                   "layers": shp["layers"], "vocab": shp["vocab"], "batch": B, "tp": n},
        "cpu_baseline": {"value": val, "unit": "tokens/s", "cores": threads * n,
                         "kind": "reference" if kind == "reference" else "port",
                         "detail": kind, "sample": sample},
        "e2e": {"value": val, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--variant", default="auto", choices=["auto", "oneshot", "twoshot"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-vllm", action="store_true", help="reference arm: skip vLLM, time the port")
    args = ap.parse_args()
    # Libraries (NCCL's version banner, vLLM's logger) write to stdout; the contract is ONE JSON
    # line there.  Everything else is sent to stderr while the run is in progress.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        res = run_reference(args) if args.impl == "reference" else run_ours(args)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    if res is not None:
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
