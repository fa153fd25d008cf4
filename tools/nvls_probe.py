#!/usr/bin/env python3
"""Where does the NVLS fused call spend its time?  Times b200c's NVLS variant at a few sizes over a
grid of role splits (copy / reduce CTAs, teams, groups) and with phases switched off through
B200C_TUNE_NVLS_DEBUG (those rows are labelled; their results are garbage, only the time counts).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29512 tools/nvls_probe.py --out gpurun_out/nvls_probe_tpN.json
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from sweep import time_graph  # noqa: E402
from vllm_distributed_b200 import _cabi as C  # noqa: E402
from vllm_distributed_b200 import comm as bc  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/nvls_probe.json")
    ap.add_argument("--hidden", type=int, default=8192)
    ap.add_argument("--tokens", type=int, nargs="+", default=[1024, 4096])
    args = ap.parse_args()
    rank, n, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    cpu = dist.new_group(backend="gloo")
    H = args.hidden
    comm = bc.B200Comm.from_process_group(cpu, dev, max_bytes=max(args.tokens) * H * 2, ag_bytes=1 << 20,
                                          p2p_bytes=1 << 20, device_group=dist.group.WORLD)
    comm.set_tuning(C.TUNE_SPIN_TIMEOUT_MS, 5000)
    assert comm.nvls, "no multicast on this box"
    gamma = torch.randn(H, device=dev).bfloat16()
    rows = []
    knobs = (C.TUNE_NVLS_COPY_BLOCKS, C.TUNE_NVLS_LINK_BLOCKS, C.TUNE_NVLS_TEAMS, C.TUNE_NVLS_GROUPS,
             C.TUNE_NVLS_DEBUG)

    def run(T, label, cpy=0, red=0, teams=0, groups=0, debug=0, plain=False):
        for k, v in zip(knobs, (cpy, red, teams, groups, debug)):
            comm.set_tuning(k, v)
        if plain:
            fn = lambda i: comm.all_reduce(xs[i % 4], C.NVLS, out)  # noqa: E731
        else:
            fn = lambda i: comm.all_reduce_add_rmsnorm_into(xs[i % 4], res, gamma, 1e-5, out, C.NVLS)  # noqa: E731
        us = time_graph(fn, 6, 3, dev, n)
        S = T * H * 2
        row = {"tokens": T, "bytes": S, "label": label, "copy": cpy, "reduce": red, "teams": teams,
               "groups": groups, "debug": debug, "plain": plain, "us": round(us, 2),
               "bus_gbs": round(2 * (n - 1) / n * S / us / 1e3, 1)}
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), file=sys.stderr, flush=True)

    for T in args.tokens:
        xs = [torch.randn(T, H, device=dev).bfloat16() for _ in range(4)]
        res = torch.randn(T, H, device=dev).bfloat16()
        out = torch.empty_like(res)
        us2 = time_graph(lambda i: comm.all_reduce_add_rmsnorm_into(xs[i % 4], res, gamma, 1e-5, out, C.TWOSHOT),
                         6, 3, dev, n)
        if rank == 0:
            print(json.dumps({"tokens": T, "label": "push two-shot (reference point)", "us": round(us2, 2)}),
                  file=sys.stderr, flush=True)
        rows.append({"tokens": T, "label": "push two-shot", "us": round(us2, 2)})
        run(T, "default")
        run(T, "default, plain all-reduce", plain=True)
        for groups in (2, 4, 8, 14):
            for teams in (1, 2, 4, 7):
                if teams <= groups:
                    run(T, f"groups {groups} teams {teams}", groups=groups, teams=teams)
        for red, teams, groups in ((24, 2, 14), (48, 2, 14), (48, 4, 14), (64, 4, 14), (56, 7, 14), (48, 8, 8)):
            run(T, f"reduce CTAs {red} teams {teams} groups {groups}", red=red, teams=teams, groups=groups)
        for dbg, what in ((1, "no copy-in"), (8, "no consumer"), (9, "no copy-in, no consumer"),
                          (4, "no multimem.st (+ consumer off)"), (2, "no ld_reduce"),
                          (6, "no ld_reduce, no multimem.st (+ consumer off)")):
            d = dbg | (8 if dbg & 4 else 0)
            run(T, f"debug: {what}", debug=d)
        comm.set_tuning(C.TUNE_NVLS_DEBUG, 0)
        # one clean call so that the buffers are consistent again before the health check
        run(T, "default (again)")
        comm.check_health()
    if rank == 0:
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        json.dump({"n": n, "hidden": H, "rows": rows}, open(args.out, "w"), indent=1)
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
