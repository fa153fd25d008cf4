#!/usr/bin/env python3
"""Per-size sweep of the fused all-reduce + residual + RMSNorm (BASELINE config 5) on real GPUs.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tools/sweep.py --out gpurun_out/sweep_tpN.json

For every (hidden, tokens) it times, as CUDA graphs of back-to-back calls (max over ranks):
  ours one-shot / two-shot (several CTA counts), the unfused pair NCCL all-reduce + our norm
  kernel, NCCL alone, and -- when it JIT-builds on the box -- FlashInfer's fused
  trtllm_allreduce_fusion (the reference's Blackwell kernel).
Prints a table on rank 0 and writes JSON; the measured crossover feeds the AUTO threshold.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# FlashInfer's fused kernel is JIT-built; a copy cross-compiled in the build container (6.5 min of
# nvcc, `gen_trtllm_comm_module().build()` with FLASHINFER_CUDA_ARCH_LIST=10.0a) travels in
# baseline/_fi/aot and is loaded as a prebuilt module, so the GPU box does not spend its lease
# compiling the comparator.
_FI_AOT = os.path.join(ROOT, "baseline", "_fi", "aot")


def _use_prebuilt_flashinfer():
    if os.path.isdir(_FI_AOT):
        import pathlib
        from flashinfer.jit import env as jit_env
        jit_env.FLASHINFER_AOT_DIR = pathlib.Path(_FI_AOT)


import contextlib  # noqa: E402

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from vllm_distributed_b200 import _cabi as C  # noqa: E402
from vllm_distributed_b200 import comm as bc  # noqa: E402


def time_graph(fn, iters, reps, dev, n, capture_ctx=None):
    """fn(i) enqueues call i; returns us per call (min over reps, max over ranks)."""
    capture_ctx = capture_ctx or contextlib.nullcontext
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(3):
            fn(i)
    torch.cuda.synchronize()
    dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s), capture_ctx():
        with torch.cuda.graph(g, stream=s):
            for i in range(iters):
                fn(i)
    torch.cuda.synchronize()
    best = None
    for _ in range(reps):
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(s):
            a.record(s)
            g.replay()
            b.record(s)
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        best = ms if best is None else min(best, ms)
    t = torch.tensor([best], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    del g
    return float(t.item()) * 1e3 / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/sweep.json")
    ap.add_argument("--hidden", type=int, nargs="+", default=[8192, 4096])
    ap.add_argument("--max-tokens", type=int, default=8192)
    ap.add_argument("--blocks", type=int, nargs="+", default=[148, 296])
    ap.add_argument("--link", type=int, nargs="+", default=[24, 48, 74, 110])
    ap.add_argument("--no-flashinfer", action="store_true")
    ap.add_argument("--no-vllm-ca", action="store_true")
    ap.add_argument("--no-symm-mem", action="store_true")
    ap.add_argument("--nvls-grid", type=str, nargs="+", default=["0:0", "24:24", "48:48", "32:64", "64:32", "74:74"],
                    help="NVLS copy:reduce CTA counts to try (0:0 = the library's default)")
    ap.add_argument("--pdl", type=int, nargs="+", default=[1, 0])
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ["RANK"])
    n = int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    cpu = dist.new_group(backend="gloo")
    max_bytes = args.max_tokens * max(args.hidden) * 2
    comm = bc.B200Comm.from_process_group(cpu, dev, max_bytes=max_bytes, ag_bytes=1 << 20,
                                          p2p_bytes=1 << 20, device_group=dist.group.WORLD)
    comm.set_tuning(C.TUNE_SPIN_TIMEOUT_MS, 5000)
    defaults = {k: comm.get_tuning(k) for k in (C.TUNE_ONESHOT_BLOCKS, C.TUNE_TWOSHOT_BLOCKS,
                                                C.TUNE_LINK_BLOCKS, C.TUNE_ONESHOT_SPLIT, C.TUNE_PDL)}
    if rank == 0:
        print(f"NVLS multicast: {'attached' if comm.nvls else 'not available'}", file=sys.stderr)

    # K7: stock vLLM's choice above 1-2 MiB on cc 10.0 -- torch symmetric memory, in-switch
    # reduction, with its copy-in / copy-out (symm_mem.py:119-154)
    k7 = None
    if not args.no_symm_mem:
        try:
            import torch.distributed._symmetric_memory as symm_mem
            k7_buf = symm_mem.empty(max_bytes // 2, dtype=torch.bfloat16, device=dev)
            k7_hdl = symm_mem.rendezvous(k7_buf, dist.group.WORLD.group_name)
            if k7_hdl.multicast_ptr:
                k7 = (k7_buf, dist.group.WORLD.group_name)
        except Exception as e:  # noqa: BLE001
            if rank == 0:
                print(f"symm_mem comparator unavailable: {type(e).__name__}: {e}", file=sys.stderr)

    fi = None
    if not args.no_flashinfer:
        try:
            _use_prebuilt_flashinfer()
            import flashinfer.comm as ficomm
            t0 = time.time()
            ws = ficomm.create_allreduce_fusion_workspace(
                backend="trtllm", world_size=n, rank=rank, max_token_num=args.max_tokens,
                hidden_dim=max(args.hidden), dtype=torch.bfloat16, group=dist.group.WORLD)
            fi = (ficomm, ws)
            if rank == 0:
                print(f"flashinfer workspace ready in {time.time() - t0:.1f}s", file=sys.stderr)
        except Exception as e:  # noqa: BLE001
            if rank == 0:
                print(f"flashinfer comparator unavailable: {type(e).__name__}: {e}", file=sys.stderr)

    ca = None
    if not args.no_vllm_ca:
        try:
            from vllm.distributed.device_communicators.custom_all_reduce import CustomAllreduce
            ca = CustomAllreduce(group=cpu, device=dev, max_size=8 << 20)
            if ca.disabled:
                ca = None
            import vllm._custom_ops as vops  # noqa: F401
        except Exception as e:  # noqa: BLE001
            ca = None
            if rank == 0:
                print(f"vLLM custom all-reduce comparator unavailable: {type(e).__name__}: {e}",
                      file=sys.stderr)
    if rank == 0:
        print(f"comparators: flashinfer={'yes' if fi else 'no'} vllm_custom_ar={'yes' if ca else 'no'}",
              file=sys.stderr)

    rows = []
    tokens = [1 << i for i in range(0, 14) if (1 << i) <= args.max_tokens]
    if args.quick:
        tokens = [t for t in tokens if t in (1, 16, 64, 256, 1024, 4096, 8192)]
    NBUF = 4
    for H in args.hidden:
        gamma = torch.randn(H, device=dev).bfloat16()
        for T in tokens:
            S = T * H * 2
            iters = 20 if S <= (4 << 20) else 6
            xs = [torch.randn(T, H, device=dev).bfloat16() for _ in range(NBUF)]
            res = torch.randn(T, H, device=dev).bfloat16()
            out = torch.empty_like(res)
            row = {"hidden": H, "tokens": T, "bytes": S, "n": n}

            def ours(variant):
                return lambda i: comm.all_reduce_add_rmsnorm_into(xs[i % NBUF], res, gamma, 1e-5,
                                                                 out, variant)
            if S * n <= 2 * max_bytes and S <= (8 << 20):
                comm.set_tuning(C.TUNE_ONESHOT_BLOCKS, 1184)
                for split in (0, 1):
                    comm.set_tuning(C.TUNE_ONESHOT_SPLIT, split)
                    for pdl in args.pdl:
                        comm.set_tuning(C.TUNE_PDL, pdl)
                        row[f"oneshot_split{split}_pdl{pdl}_us"] = round(
                            time_graph(ours(C.ONESHOT), iters, 3, dev, n), 2)
                comm.set_tuning(C.TUNE_PDL, 1)
                comm.set_tuning(C.TUNE_ONESHOT_SPLIT, 0)
                row["oneshot_us"] = min(v for k, v in row.items() if k.startswith("oneshot_split"))
            best = None
            for blk in args.blocks:
                for link in args.link:
                    if link >= blk:
                        continue
                    comm.set_tuning(C.TUNE_TWOSHOT_BLOCKS, blk)
                    comm.set_tuning(C.TUNE_LINK_BLOCKS, link)
                    us = time_graph(ours(C.TWOSHOT), iters, 3, dev, n)
                    row[f"twoshot_b{blk}_l{link}_us"] = round(us, 2)
                    best = us if best is None else min(best, us)
            comm.set_tuning(C.TUNE_LINK_BLOCKS, 0)
            row["twoshot_us"] = round(best, 2)
            # what AUTO runs with nothing tuned: the number a user of the library gets
            for k, v in defaults.items():
                comm.set_tuning(k, v)
            row["auto_us"] = round(time_graph(ours(C.AUTO), iters, 3, dev, n), 2)
            if comm.nvls:
                best_nv = None
                for spec in args.nvls_grid:
                    cpy, red = (int(v) for v in spec.split(":"))
                    comm.set_tuning(C.TUNE_NVLS_COPY_BLOCKS, cpy)
                    comm.set_tuning(C.TUNE_NVLS_LINK_BLOCKS, red)
                    us = time_graph(ours(C.NVLS), iters, 3, dev, n)
                    row[f"nvls_c{cpy}_r{red}_us"] = round(us, 2)
                    best_nv = us if best_nv is None else min(best_nv, us)
                comm.set_tuning(C.TUNE_NVLS_COPY_BLOCKS, 0)
                comm.set_tuning(C.TUNE_NVLS_LINK_BLOCKS, 0)
                row["nvls_us"] = round(best_nv, 2)
                row["ar_nvls_us"] = round(time_graph(
                    lambda i: comm.all_reduce(xs[i % NBUF], C.NVLS, out), iters, 3, dev, n), 2)
            # plain all-reduce (no epilogue), both variants at default CTA counts
            comm.set_tuning(C.TUNE_ONESHOT_BLOCKS, 148)
            comm.set_tuning(C.TUNE_TWOSHOT_BLOCKS, 148)
            if S * n <= 2 * max_bytes and S <= (8 << 20):
                row["ar_oneshot_us"] = round(time_graph(
                    lambda i: comm.all_reduce(xs[i % NBUF], C.ONESHOT, out), iters, 3, dev, n), 2)
            row["ar_twoshot_us"] = round(time_graph(
                lambda i: comm.all_reduce(xs[i % NBUF], C.TWOSHOT, out), iters, 3, dev, n), 2)

            # NCCL all-reduce alone, and the unfused pair (NCCL + our local norm kernel)
            tmp = [torch.empty_like(x) for x in xs]

            def nccl_only(i):
                tmp[i % NBUF].copy_(xs[i % NBUF])
                dist.all_reduce(tmp[i % NBUF])

            def nccl_pair(i):
                nccl_only(i)
                bc.fused_add_rms_norm_(tmp[i % NBUF], res, gamma, 1e-5)
            row["nccl_ar_us"] = round(time_graph(nccl_only, iters, 3, dev, n), 2)
            row["nccl_pair_us"] = round(time_graph(nccl_pair, iters, 3, dev, n), 2)

            if k7 is not None:
                try:
                    kbuf, gname = k7
                    view = kbuf[: T * H]
                    o7 = torch.empty(T, H, dtype=torch.bfloat16, device=dev)

                    def k7_only(i):
                        view.copy_(xs[i % NBUF].view(-1))
                        torch.ops.symm_mem.multimem_all_reduce_(view, "sum", gname)
                        o7.copy_(view.view(T, H))

                    def k7_pair(i):
                        k7_only(i)
                        bc.fused_add_rms_norm_(o7, res, gamma, 1e-5)
                    row["symm_mem_ar_us"] = round(time_graph(k7_only, iters, 3, dev, n), 2)
                    row["symm_mem_pair_us"] = round(time_graph(k7_pair, iters, 3, dev, n), 2)
                except Exception as e:  # noqa: BLE001
                    row["symm_mem_error"] = f"{type(e).__name__}: {e}"[:200]

            if ca is not None and S <= (8 << 20):
                try:
                    def ca_only(i):
                        return ca.custom_all_reduce(xs[i % NBUF])

                    def ca_pair(i):
                        o = ca.custom_all_reduce(xs[i % NBUF])
                        torch.ops._C.fused_add_rms_norm(o, res, gamma, 1e-5)
                    row["vllm_ca_us"] = round(time_graph(ca_only, iters, 3, dev, n, ca.capture), 2)
                    row["vllm_ca_pair_us"] = round(time_graph(ca_pair, iters, 3, dev, n, ca.capture), 2)
                except Exception as e:  # noqa: BLE001
                    row["vllm_ca_error"] = f"{type(e).__name__}: {e}"[:200]

            if fi is not None:
                ficomm, ws = fi
                try:
                    ro, no = torch.empty_like(res), torch.empty_like(res)

                    def fi_call(i, oneshot=None):
                        ficomm.allreduce_fusion(
                            input=xs[i % NBUF], workspace=ws,
                            pattern=ficomm.AllReduceFusionPattern.kARResidualRMSNorm,
                            launch_with_pdl=True, trigger_completion_at_end=T > 16,
                            residual_out=ro, norm_out=no, residual_in=res, rms_gamma=gamma,
                            rms_eps=1e-5, use_oneshot=(True if T <= n else oneshot), fp32_acc=True)
                    row["flashinfer_auto_us"] = round(time_graph(fi_call, iters, 3, dev, n), 2)
                except Exception as e:  # noqa: BLE001
                    row["flashinfer_error"] = f"{type(e).__name__}: {e}"[:200]
            bus = 2 * (n - 1) / n * S
            best = min(v for k, v in row.items() if k in ("oneshot_us", "twoshot_us", "nvls_us"))
            row["best_us"] = best
            row["bus_gbs"] = round(bus / best / 1e3, 1)
            rows.append(row)
            if rank == 0:
                print(json.dumps(row), file=sys.stderr, flush=True)
            comm.check_health()
    if rank == 0:
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        json.dump({"n": n, "rows": rows}, open(args.out, "w"), indent=1)
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
