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


_REAL_STDOUT_FD = None  # set by main() while fd 1 is pointed at stderr


def real_print(line: str) -> None:
    """Write one line to the process's real stdout (the JSON contract), even mid-run."""
    if _REAL_STDOUT_FD is None:
        print(line, flush=True)
    else:
        os.write(_REAL_STDOUT_FD, (line + "\n").encode())


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


def workload_config(shp: dict, n: int, B: int, variant: str) -> dict:
    """The `config` object of the JSON line; both arms print the same one for the same flags."""
    calls = 2 * shp["layers"] + 1
    return {
        "workload": f"{shp['model']} TP={n} decode hot path, batch {B}: {calls} x "
                    + ("fused residual-add+RMSNorm" if n == 1 else
                       "fused all-reduce+residual-add+RMSNorm + 1 lm_head all-gather")
                    + f" on [{B},{shp['hidden']}] bf16",
        "hidden": shp["hidden"], "layers": shp["layers"], "vocab": shp["vocab"], "batch": B, "tp": n,
        "calls_per_step": calls + (1 if n > 1 else 0),
        "l2": "256 MiB buffer rewritten between timed steps (untimed) to flush the 126 MB L2",
        "execution": "step captured once as a CUDA graph, replayed per timed step",
        "variant": variant,
    }


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
    extras = n > 1 and not args.no_extras
    big_tokens = 4096  # the bandwidth-regime point of the north star: [4096, H] = 64 MiB at H = 8192
    if n > 1:
        dist.init_process_group(backend="nccl", device_id=dev)
        cpu_group = dist.new_group(backend="gloo")
        max_bytes = max(S, 1 << 20, big_tokens * H * 2 if extras else 0)
        comm = bc.B200Comm.from_process_group(cpu_group, dev, max_bytes=max_bytes,
                                              ag_bytes=max(B * V * 2, 1 << 20), p2p_bytes=1 << 20,
                                              device_group=dist.group.WORLD)
        log(f"[rank {rank}] arenas up, NVLS multicast {'attached' if comm.nvls else 'not available'}")
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
    variant = {"auto": C.AUTO, "oneshot": C.ONESHOT, "twoshot": C.TWOSHOT, "nvls": C.NVLS}[args.variant]

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

    # ---- parity of the REAL multi-GPU kernels, checked where the driver can see it ------------------
    # One eager step against NCCL + an fp32 torch norm (vLLM's native definition,
    # ir/ops/layernorm.py:39-58).  Two references for the all-reduce:
    #   "f32"  NCCL sum of the fp32-upcast inputs, rounded once -- the reference custom all-reduce's
    #          arithmetic (fp32 accumulation, one rounding, vllm_custom_all_reduce.cuh:213-221) and
    #          this library's; only the fp32 summation order can differ.  Used for the first call
    #          AND for the whole chained step (161 dependent calls at TP=8).
    #   "bf16" NCCL sum of the bf16 inputs as stock vLLM's NCCL path computes it (it rounds
    #          partial sums): north_star's per-layer bound, first call only -- chained over 161
    #          dependent layers two valid roundings drift apart by more than any per-layer bound.
    # Tolerance 1e-2 relative (Frobenius) everywhere; at N=2 the residual of the first call must
    # be bit-exact (one add, one rounding on every path); the logits gather bit-exact.
    # A miss aborts the run with a non-zero exit.
    def rel(a, b):
        return float((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-30))

    def ref_call(xi, res, g, f32=True):
        if f32:
            s32 = xi.float()
            dist.all_reduce(s32)
            sref = s32.to(torch.bfloat16)
        else:
            sref = xi.clone()
            dist.all_reduce(sref)
        x32 = sref.float() + res.float()
        rnew = x32.to(torch.bfloat16)
        y = (x32 * torch.rsqrt(x32.pow(2).mean(-1, keepdim=True) + eps)).to(torch.bfloat16) * g
        return y, rnew

    parity = None
    if n > 1:
        checks, worst = {}, 0.0
        variants = [("auto", variant)]
        if comm.nvls and variant != C.NVLS:
            variants.append(("nvls", C.NVLS))
        for vname, vcode in variants:
            residual.copy_(res_host)
            comm.all_reduce_add_rmsnorm_into(x[0], residual, gammas[0], eps, h, vcode)
            ok = True
            for f32 in (True, False):
                y_ref, r_ref = ref_call(x[0], res_host.to(dev), gammas[0], f32)
                torch.cuda.synchronize()
                e_r, e_y = rel(residual, r_ref), rel(h, y_ref)
                ok = ok and e_r <= 1e-2 and e_y <= 1e-2
                if n == 2 and vcode != C.NVLS:
                    ok = ok and torch.equal(residual, r_ref)
                worst = max(worst, e_r, e_y)
            checks[f"first_call_{vname}"] = ok
        # the whole step, chained through the residual stream, exactly as it is timed
        residual.copy_(res_host)
        step()
        r_ref = res_host.to(dev)
        for i in range(calls):
            y_ref, r_ref = ref_call(x[i], r_ref, gammas[i])
        gathered = torch.empty(n, B, V // n, dtype=torch.bfloat16, device=dev)
        dist.all_gather_into_tensor(gathered, logits_local)
        logits_ref = gathered.permute(1, 0, 2).reshape(B, V)
        torch.cuda.synchronize()
        e_r, e_y = rel(residual, r_ref), rel(h, y_ref)
        checks["step_residual"] = e_r <= 1e-2
        checks["step_norm_out"] = e_y <= 1e-2
        checks["logits_gather_bit_exact"] = bool(torch.equal(logits_full, logits_ref))
        worst = max(worst, e_r, e_y)
        # Informational, never gating: the reference's own custom all-reduce kernels (K1/K2,
        # fp32 accumulation in rank order, one rounding) on the same input must give the SAME BITS
        # as this library's all-reduce -- the GPU-box oracle of SURVEY 8c.
        code, why = 0, "unavailable"   # 0 unavailable, 1 differs, 2 same bits
        try:
            from vllm.distributed.device_communicators.custom_all_reduce import CustomAllreduce
            ca = CustomAllreduce(group=cpu_group, device=dev, max_size=8 << 20)
            if ca.disabled:
                why = "unavailable: vLLM custom all-reduce disabled itself on this box"
            else:
                xs_half = x[0][: max(1, B // 2)].contiguous()  # under the reference's 1 MiB cap at TP=8
                want = ca.custom_all_reduce(xs_half)
                got = comm.all_reduce(xs_half)
                torch.cuda.synchronize()
                if want is None:
                    why = "unavailable: the reference kernel does not take this size"
                else:
                    code = 2 if torch.equal(got, want) else 1
                ca.close()
        except Exception as e:  # noqa: BLE001 -- a comparator problem must not fail the bench
            why = f"unavailable: {type(e).__name__}: {str(e)[:120]}"
        st = torch.tensor([code], dtype=torch.int32, device=dev)
        dist.all_reduce(st, op=dist.ReduceOp.MIN)   # the one collective of this block, on every path
        bitwise = {2: True, 1: False}.get(int(st.item()), why)
        flag = torch.tensor([0 if all(checks.values()) else 1], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MAX)
        w = torch.tensor([worst], dtype=torch.float64, device=dev)
        dist.all_reduce(w, op=dist.ReduceOp.MAX)
        parity = {"ok": int(flag.item()) == 0, "max_rel": round(float(w.item()), 6),
                  "checks_rank0": checks, "bitwise_vs_vllm_custom_all_reduce": bitwise,
                  "against": "NCCL all_reduce (fp32-upcast sum rounded once; and, first call, the bf16 sum) / "
                             "all_gather_into_tensor + fp32 torch norm, every rank"}
        if not parity["ok"]:
            log(f"[rank {rank}] PARITY FAILURE: {checks} max_rel {worst}")
            dist.barrier()
            if rank == 0:
                real_print(json.dumps({"metric": METRIC, "n_gpus": n, "parity": parity,
                                       "error": "parity check failed; nothing was timed"}))
            raise SystemExit(3)
        residual.copy_(res_host)

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

    # ---- N>1: the numbers the north star is stated in ----------------------------------------------
    # (a) the step split into its fused all-reduce calls and the logits gather, each as its own
    # graph; (b) one fused call per regime -- [64, H] (latency) and [4096, H] (bandwidth) -- for
    # every variant that applies, as bus GB/s = 2(N-1)/N * S / t against the measured link peak.
    per_call_us = gather_us = None
    regimes = []
    if n > 1:
        def graph_of(fn):
            with torch.cuda.stream(stream):
                fn()
            sync_all()
            gph = torch.cuda.CUDAGraph()
            with torch.cuda.stream(stream):
                with torch.cuda.graph(gph, stream=stream):
                    fn()
            sync_all()
            return gph

        def ar_only():
            for i in range(calls):
                comm.all_reduce_add_rmsnorm_into(x[i], residual, gammas[i], eps, h, variant)

        k2 = max(5, min(args.steps, 20))
        ms_ar, _ = timed_steps(graph_of(ar_only).replay, k2, 3)
        ms_ag, _ = timed_steps(graph_of(lambda: comm.all_gather_lastdim(logits_local, out=logits_full)).replay, k2, 3)
        per_call_us = round(ms_ar * 1e3 / calls, 3)
        gather_us = round(ms_ag * 1e3, 3)
    if extras:
        for tokens in (B, big_tokens):
            # enough calls per graph that the graph's own launch cost (~10 us) disappears
            reps = 8 if tokens * H * 2 > (4 << 20) else 64
            xb = torch.randn(tokens, H, generator=gen).bfloat16().to(dev)
            rb = torch.randn(tokens, H, generator=g7).bfloat16().to(dev)
            hb = torch.empty_like(xb)
            Sb = tokens * H * 2
            names = [("auto", C.AUTO), ("twoshot", C.TWOSHOT)]
            if Sb * n <= 2 * comm.max_bytes and Sb <= (4 << 20):
                names.append(("oneshot", C.ONESHOT))
            if comm.nvls:
                names.append(("nvls", C.NVLS))
            for vname, vcode in names:
                def rep(vcode=vcode):
                    for _ in range(reps):
                        comm.all_reduce_add_rmsnorm_into(xb, rb, gammas[0], eps, hb, vcode)
                ms, _ = timed_steps(graph_of(rep).replay, 5, 3)
                us = ms * 1e3 / reps
                bus = 2 * (n - 1) / n * Sb / (us * 1e-6) / 1e9
                regimes.append({"shape": [tokens, H], "bytes": Sb, "variant": vname, "us_per_call": round(us, 3),
                                "bus_gbs": round(bus, 1), "frac": round(bus / NVLINK_PEAK_GBS, 4)})
            del xb, rb, hb

    # ---- N=1: what would a call cost if it moved (almost) nothing? ----------------------------------
    # The batch-64 call is latency-bound (one DRAM round trip + one block reduction per 2 MB), so
    # the HBM fraction says little; the same 65 launches on ONE row are the floor of this launch
    # chain (PDL-linked kernels in one graph), and the step is reported as a multiple of it.
    floor_us = None
    if n == 1:
        x1 = torch.randn(calls, 1, H, generator=gen).bfloat16().to(dev)
        r1 = torch.randn(1, H, generator=g7).bfloat16().to(dev)
        h1 = torch.empty(1, H, dtype=torch.bfloat16, device=dev)

        def floor_step():
            for i in range(calls):
                comm.all_reduce_add_rmsnorm_into(x1[i], r1, gammas[i], eps, h1)
        with torch.cuda.stream(stream):
            floor_step()
        sync_all()
        gfloor = torch.cuda.CUDAGraph()
        with torch.cuda.stream(stream):
            with torch.cuda.graph(gfloor, stream=stream):
                floor_step()
        sync_all()
        ms_floor, _ = timed_steps(gfloor.replay, max(5, min(args.steps, 20)), 3)
        floor_us = ms_floor * 1e3 / calls

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
                "algorithmic_bytes_per_launch": alg_bytes,
                "launch_floor_us": round(floor_us, 3),
                "call_over_floor": round(ms_per_step * 1e3 / calls / floor_us, 3),
                "note": "latency-bound at batch 64: launch_floor_us = the same graph of 65 PDL-linked "
                        "launches on ONE row; bandwidth-regime numbers for this kernel are in "
                        "profiles/r2_norm_sweep_tp1.json (1.0 of the measured HBM copy peak at 16384 rows)"}
    else:
        bus_bytes = 2 * (n - 1) / n * S  # all-reduce bus bytes
        achieved = bus_bytes * calls / (ms_per_step * 1e-3) / 1e9
        if per_call_us:  # the fused calls alone (the gather is timed separately)
            achieved = bus_bytes / (per_call_us * 1e-6) / 1e9
        roof = {"bound": "nvlink", "kernel": "ar_norm_kernel<bf16> / ar_norm_cluster_kernel<bf16> / nvls_kernel<bf16> (by size)",
                "achieved": round(achieved, 2),
                "peak": NVLINK_PEAK_GBS, "unit": "GB/s", "frac": round(achieved / NVLINK_PEAK_GBS, 4),
                "traffic": None, "per_call_us": per_call_us, "gather_us": gather_us, "regimes": regimes,
                "peak_source": "measured peer copy 770 GB/s per direction (B200_PROFILING.md); "
                               "nominal 900",
                "algorithmic_bytes_per_launch": bus_bytes,
                "note": "achieved = bus bytes of one fused call / its time inside a graph of the step's "
                        "fused calls alone; `regimes` = one fused call per size and variant, same process; "
                        "with NVLS every byte crosses the link once, so bus GB/s (ring convention) can "
                        "exceed the per-direction link peak"}

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
        "scaling_note": "every rank's message (batch x hidden) is the same size at every N, so per-GPU "
                        "work on this path is fixed -- but the per-N lines are different model shapes "
                        "(8B at N<=2, 70B at N>=4), so efficiency across N compares different "
                        "workloads; track us per fused call instead (DESIGN.md 6)",
        "vs_baseline": None,
        "dtype": "bf16",
        "data": "synthetic",
        "parity": parity,
        "nvls": bool(getattr(comm, "nvls", False)),
        "config": workload_config(shp, n, B, args.variant),
        "clocks": clocks,
        "e2e": {"value": round(B / (e2e_ms * 1e-3), 1), "unit": "tokens/s",
                "ms_per_step": round(e2e_ms, 5), "steps": e2e_steps,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "path": "B200Comm python API, eager launches, pinned host buffers"},
        "gpu_launches": launches_per_step * args.steps,
        "roofline": roof,
        "us_per_call": round(ms_per_step * 1e3 / launches_per_step, 3),
    }
    if rank == 0 and not args.no_cpu_baseline:
        result["cpu_baseline"] = cpu_baseline_port(shp, B, calls, n, budget_s=10.0 if n == 1 else 5.0)
    if n > 1:
        dist.barrier()
        comm.close()
        dist.destroy_process_group()
    return result if rank == 0 else None


# ================================ CPU baselines ==========================================================

def cpu_baseline_port(shp: dict, B: int, calls: int, n: int = 1, budget_s: float = 10.0) -> dict:
    """The oracle (a port of the reference's algorithm) timed on the host cores: the same
    step (all N ranks' all-reduce + residual + RMSNorm in one process at N > 1), as many
    repetitions as fit in ~budget_s."""
    import torch
    from oracle import oracle
    H, eps = shp["hidden"], shp["eps"]
    cores = os.cpu_count() or 1
    oracle.set_num_threads(cores)
    g = torch.Generator().manual_seed(1)
    xs = [[torch.randn(B, H, generator=g).bfloat16() for _ in range(n)] for _ in range(4)]
    res = torch.randn(B, H, generator=g).bfloat16()
    gamma = torch.randn(H, generator=g).bfloat16()
    t0 = time.perf_counter()
    steps = 0
    while True:
        for i in range(calls):
            if n == 1:
                _, res = oracle.fused_add_rms_norm(xs[i % 4][0], res, gamma, eps)
            else:
                _, res, _ = oracle.all_reduce_add_rms_norm(xs[i % 4], res, gamma, eps)
        steps += 1
        if time.perf_counter() - t0 > budget_s or steps >= 200:
            break
    dt = (time.perf_counter() - t0) / steps
    what = "fused residual-add+RMSNorm" if n == 1 else f"{n}-rank all-reduce + residual-add+RMSNorm"
    return {"value": round(B / dt, 1), "unit": "tokens/s", "cores": oracle.num_threads(),
            "kind": "port", "ms_per_step": round(dt * 1e3, 3),
            "sample": f"{steps} full steps ({calls} {what} calls on [{B},{H}] bf16 each"
                      + ("" if n == 1 else ", logits gather not included")
                      + ") of oracle/oracle.c with OpenMP"}


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
        # This arm times vLLM's CPU (gloo) communicator path.  On a GPU box vLLM's platform
        # detection (NVML sees GPUs even with CUDA_VISIBLE_DEVICES="") would pick CudaPlatform
        # and then fail to build its NCCL communicator; switch the CUDA detector off so that the
        # platform resolves exactly as on a CPU-only host (where the golden fixtures were made).
        import vllm.platforms as vllm_platforms
        vllm_platforms.builtin_platform_plugins["cuda"] = lambda: None
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
        kind = f"port ({type(e).__name__}: {str(e)[:160]})"
        log(f"[reference arm rank {rank}] vLLM CPU path unusable, timing the port instead: {e!r}")
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


# Variables torch.distributed.run exports into its ranks.  The reference arm starts its OWN
# gloo group from rank 0, so none of them may leak into the spawned workers: with
# TORCHELASTIC_USE_AGENT_STORE=True every rank of a tcp:// rendezvous becomes a TCPStore
# client and nobody serves (the round-1 hang at N>1).
_TORCHRUN_ENV_PREFIXES = ("TORCHELASTIC_", "GROUP_", "ROLE_", "MASTER_", "TORCH_NCCL_", "PET_")
_TORCHRUN_ENV_NAMES = ("RANK", "LOCAL_RANK", "WORLD_SIZE", "LOCAL_WORLD_SIZE", "NODE_RANK",
                       "OMP_NUM_THREADS")


def _scrub_torchrun_env() -> None:
    for k in list(os.environ):
        if k in _TORCHRUN_ENV_NAMES or k.startswith(_TORCHRUN_ENV_PREFIXES):
            del os.environ[k]


def _free_port() -> int:
    import socket
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        return sk.getsockname()[1]


def run_reference(args) -> dict | None:
    """--impl reference: the reference's own CPU implementation of the path (vLLM's
    communicator API over gloo + its native fused_add_rms_norm) on the box's host cores.
    Under torchrun only rank 0 works; it spawns its own N gloo workers."""
    if int(os.environ.get("RANK", 0)) != 0:
        return None
    _scrub_torchrun_env()
    import queue as queue_mod
    import torch.multiprocessing as mp
    n = args.gpus
    shp = model_shape(n)
    B, calls = args.batch, 2 * shp["layers"] + 1
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_ref_worker,
                         args=(r, n, port, shp, B, calls, steps, warmup, not args.no_vllm, q))
             for r in range(n)]
    for p in procs:
        p.start()
    outs, deadline = [], time.time() + args.ref_timeout
    while len(outs) < n:
        try:
            outs.append(q.get(timeout=1.0))
            continue
        except queue_mod.Empty:
            pass
        dead = [p for p in procs if p.exitcode not in (None, 0)]
        if dead or time.time() > deadline:
            for p in procs:
                if p.is_alive():
                    p.terminate()
            why = (f"worker exited with code {dead[0].exitcode}" if dead
                   else f"no result within {args.ref_timeout} s")
            raise SystemExit(f"reference arm failed: {why}")
    for p in procs:
        p.join(timeout=60)
    dt = max(o[1] for o in outs)
    kind = outs[0][2]
    threads = outs[0][3]
    val = round(B / dt, 1)
    sample = (f"{steps} full steps (after {warmup} warm-up) of the same workload, {n} gloo "
              f"process(es) x {threads} threads")
    return {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "tokens/s", "n_gpus": n,
        "steps": steps, "warmup": warmup, "ms_per_step": round(dt * 1e3, 3),
        "higher_is_better": True, "scaling": "weak",
        "scaling_note": "every rank's message (batch x hidden) is the same size at every N, so per-GPU "
                        "work on this path is fixed -- but the per-N lines are different model shapes "
                        "(8B at N<=2, 70B at N>=4), so efficiency across N compares different "
                        "workloads; track us per fused call instead (DESIGN.md 6)",
        "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": workload_config(shp, n, B, args.variant),
        "cpu_baseline": {"value": val, "unit": "tokens/s", "cores": threads * n,
                         "kind": "reference" if kind == "reference" else "port",
                         "detail": kind + ": vLLM CPU/gloo communicator path on the host cores",
                         "sample": sample},
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
    ap.add_argument("--variant", default="auto", choices=["auto", "oneshot", "twoshot", "nvls"])
    ap.add_argument("--no-extras", action="store_true",
                    help="N>1: skip the per-regime table (1 MiB and 64 MiB fused calls, every variant)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-vllm", action="store_true", help="reference arm: skip vLLM, time the port")
    ap.add_argument("--ref-timeout", type=float, default=1200.0,
                    help="reference arm: give up if the gloo workers have not reported by then")
    args = ap.parse_args()
    # Libraries (NCCL's version banner, vLLM's logger) write to stdout; the contract is ONE JSON
    # line there.  Everything else is sent to stderr while the run is in progress.
    global _REAL_STDOUT_FD
    sys.stdout.flush()
    real_stdout = os.dup(1)
    _REAL_STDOUT_FD = real_stdout
    os.dup2(2, 1)
    try:
        res = run_reference(args) if args.impl == "reference" else run_ours(args)
    finally:
        sys.stdout.flush()
        _REAL_STDOUT_FD = None
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    if res is not None:
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
