#!/usr/bin/env python3
"""Pipeline-stage hop microbenchmark (PP bytes = 2*S per hop), 2 ranks:

  * device side: {"hidden_states", "residual"} of S bytes each, rank 0 -> rank 1 and back
    (ping-pong, eager launches, CUDA events; one-way time = round trip / 2), through
    B200Comm.send/recv (mailbox kernels) and through NCCL isend/irecv as vLLM's
    send/recv_tensor_dict issues them (parallel_state.py:871-926, 966-1050);
  * host side: the per-hop metadata message, pickled over gloo exactly as
    GroupCoordinator.send_object does (two blocking sends, parallel_state.py:674-735) against
    pp_channel's shared-memory ring, with fresh metadata every step and with repeated metadata
    (the "same as before" marker).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29513 tools/pp_bench.py --out gpurun_out/pp_bench.json
"""
from __future__ import annotations

import argparse
import json
import os
import pickle
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


class MiniGroup:
    """The three things pp_channel.send_meta / recv_meta use of a GroupCoordinator."""

    def __init__(self, cpu_group, rank):
        self.cpu_group, self.rank_in_group, self.ranks = cpu_group, rank, [0, 1]

    def send_object(self, obj, dst):  # as parallel_state.py:674-700
        t = torch.frombuffer(bytearray(pickle.dumps(obj)), dtype=torch.uint8)
        dist.send(torch.tensor([t.numel()], dtype=torch.long), dst=dst, group=self.cpu_group)
        dist.send(t, dst=dst, group=self.cpu_group)

    def recv_object(self, src):  # as parallel_state.py:702-735
        n = torch.empty(1, dtype=torch.long)
        dist.recv(n, src=src, group=self.cpu_group)
        t = torch.empty(int(n.item()), dtype=torch.uint8)
        dist.recv(t, src=src, group=self.cpu_group)
        return pickle.loads(t.numpy().tobytes())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/pp_bench.json")
    ap.add_argument("--cpu-only", action="store_true", help="metadata leg only (no GPU needed)")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    assert world == 2
    gpu = torch.cuda.is_available() and not args.cpu_only
    if gpu:
        torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
        dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
        dist.init_process_group("nccl", device_id=dev)
        cpu = dist.new_group(backend="gloo")
    else:
        dist.init_process_group("gloo")
        cpu = dist.group.WORLD
    peer = 1 - rank
    result = {"device": [], "metadata": {}}

    # ---- host side: metadata message ----------------------------------------------------------------
    from vllm_distributed_b200 import pp_channel
    try:
        from vllm.distributed.parallel_state import TensorMetadata
    except Exception:  # noqa: BLE001
        from collections import namedtuple
        TensorMetadata = namedtuple("TensorMetadata", ["device", "dtype", "size"])
    g = MiniGroup(cpu, rank)

    def meta(step):
        return [("hidden_states", TensorMetadata("cuda", torch.bfloat16, torch.Size([64, 8192]))),
                ("residual", TensorMetadata("cuda", torch.bfloat16, torch.Size([64, 8192]))), ("step", step)]

    def ping_pong(send, recv, n, varying):
        dist.barrier(group=cpu)
        t0 = time.perf_counter()
        for i in range(n):
            m = meta(i if varying else 0)
            if rank == 0:
                send(m)
                recv()
            else:
                recv()
                send(m)
        return (time.perf_counter() - t0) / n / 2 * 1e6  # one-way us

    n = 300
    gloo_us = ping_pong(lambda m: g.send_object(m, peer), lambda: g.recv_object(peer), n, True)
    pp_channel.send_meta(g, meta(-1), peer) if rank == 0 else pp_channel.recv_meta(g, peer)   # sets up 0 -> 1
    pp_channel.recv_meta(g, peer) if rank == 0 else pp_channel.send_meta(g, meta(-1), peer)   # sets up 1 -> 0
    shm_us = ping_pong(lambda m: pp_channel.send_meta(g, m, peer), lambda: pp_channel.recv_meta(g, peer), n, True)
    same_us = ping_pong(lambda m: pp_channel.send_meta(g, m, peer), lambda: pp_channel.recv_meta(g, peer), n, False)
    result["metadata"] = {"gloo_send_object_us": round(gloo_us, 1), "shm_ring_us": round(shm_us, 1),
                          "shm_ring_same_as_before_us": round(same_us, 1), "messages": n,
                          "note": "one-way host time per metadata message (ping-pong / 2)"}

    # ---- device side: the two tensors of a hop -----------------------------------------------------
    if gpu:
        from vllm_distributed_b200 import _cabi as C
        from vllm_distributed_b200 import comm as bc
        comm = bc.B200Comm.from_process_group(cpu, dev, max_bytes=1 << 20, ag_bytes=1 << 20, p2p_bytes=16 << 20)
        comm.set_tuning(C.TUNE_SPIN_TIMEOUT_MS, 5000)
        for S in (16 << 10, 128 << 10, 1 << 20, 4 << 20, 16 << 20):
            hs = torch.randn(S // 2, device=dev).bfloat16()
            rs = torch.randn(S // 2, device=dev).bfloat16()
            hs2, rs2 = torch.empty_like(hs), torch.empty_like(rs)

            def ours():
                if rank == 0:
                    comm.send(hs, peer); comm.send(rs, peer); comm.recv(hs2, peer); comm.recv(rs2, peer)
                else:
                    comm.recv(hs2, peer); comm.recv(rs2, peer); comm.send(hs, peer); comm.send(rs, peer)

            def nccl():
                if rank == 0:
                    for w in [dist.isend(hs, peer), dist.isend(rs, peer)]:
                        w.wait()
                    for w in [dist.irecv(hs2, peer), dist.irecv(rs2, peer)]:
                        w.wait()
                else:
                    for w in [dist.irecv(hs2, peer), dist.irecv(rs2, peer)]:
                        w.wait()
                    for w in [dist.isend(hs, peer), dist.isend(rs, peer)]:
                        w.wait()

            row = {"bytes_per_tensor": S, "hop_bytes": 2 * S}
            for name, fn in (("b200c", ours), ("nccl", nccl)):
                for _ in range(5):
                    fn()
                torch.cuda.synchronize()
                dist.barrier(group=cpu)
                iters = 30
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(iters):
                    fn()
                b.record()
                torch.cuda.synchronize()
                us = a.elapsed_time(b) * 1e3 / iters / 2
                t = torch.tensor([us], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                row[f"{name}_one_way_us"] = round(float(t.item()), 2)
                row[f"{name}_gbs"] = round(2 * S / float(t.item()) / 1e3, 1)
            result["device"].append(row)
            if rank == 0:
                print(json.dumps(row), file=sys.stderr, flush=True)
        comm.check_health()
        dist.barrier()
        comm.close()
    if rank == 0:
        print(json.dumps(result["metadata"]), file=sys.stderr, flush=True)
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        json.dump(result, open(args.out, "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
