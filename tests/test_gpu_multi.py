"""Real multi-GPU parity (one process per GPU, CUDA IPC arenas over NVLink).  Skipped when the
box has a single GPU -- the emulated-rank tests in test_gpu_parity.py cover the device code."""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from oracle import oracle
    from vllm_distributed_b200 import _cabi as C
    from vllm_distributed_b200 import comm as bc
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    cpu_group = dist.new_group(backend="gloo")
    comm = bc.B200Comm.from_process_group(cpu_group, dev, max_bytes=8 << 20,
                                          ag_bytes=8 << 20, p2p_bytes=1 << 20,
                                          device_group=dist.group.WORLD)
    comm.set_tuning(C.TUNE_SPIN_TIMEOUT_MS, 5000)
    ok = True
    notes = []
    try:
        for i, (t, h) in enumerate([(1, 896), (5, 4096), (64, 8192), (300, 4096), (1000, 4096)]):
            xs = [torch.randn(t, h, generator=torch.Generator().manual_seed(1234 + r)).bfloat16()
                  for r in range(world)]
            res = torch.randn(t, h, generator=torch.Generator().manual_seed(7)).bfloat16()
            gamma = torch.randn(h, generator=torch.Generator().manual_seed(8)).bfloat16()
            y_ref, r_ref, ar_ref = oracle.all_reduce_add_rms_norm(xs, res, gamma, 1e-5)
            for variant in (C.ONESHOT, C.TWOSHOT, C.AUTO):
                ar = comm.all_reduce(xs[rank].to(dev), variant)
                y, r = comm.all_reduce_add_rmsnorm(xs[rank].to(dev), res.to(dev), gamma.to(dev),
                                                   1e-5, variant)
                torch.cuda.synchronize()
                ok &= torch.equal(ar.cpu(), ar_ref)
                ok &= torch.equal(r.cpu(), r_ref)
                ok &= bool(torch.allclose(y.cpu().float(), y_ref.float(), atol=1e-3, rtol=1.6e-2))
            if comm.nvls:
                # in-switch reduction: the summation order is the switch's, so the bar is
                # north_star's 1e-2 relative -- and identical bits on every rank
                for groups in (0, 1, 3):
                    comm.set_tuning(C.TUNE_NVLS_GROUPS, groups)
                    ar = comm.all_reduce(xs[rank].to(dev), C.NVLS)
                    y, r = comm.all_reduce_add_rmsnorm(xs[rank].to(dev), res.to(dev), gamma.to(dev),
                                                       1e-5, C.NVLS)
                    torch.cuda.synchronize()
                    for got, ref in ((ar, ar_ref), (r, r_ref), (y, y_ref)):
                        err = float((got.cpu().float() - ref.float()).norm() / ref.float().norm())
                        ok &= err <= 1e-2
                        ok &= bool(torch.allclose(got.cpu().float(), ref.float(), atol=3e-2, rtol=3e-2))
                    mism = int((ar.cpu().view(torch.int16) != ar_ref.view(torch.int16)).sum())
                    notes.append(f"nvls [{t},{h}] g{groups}: {mism}/{ar_ref.numel()} all-reduce elements differ from the rank-order sum")
                    # every rank must hold the same bits
                    mine = ar.view(torch.int16).to(torch.int32).sum(dtype=torch.int64).reshape(1)
                    lo, hi = mine.clone(), mine.clone()
                    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
                    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
                    ok &= bool((lo == hi).item())
                comm.set_tuning(C.TUNE_NVLS_GROUPS, 0)
        slabs = [torch.randn(7, 1000, generator=torch.Generator().manual_seed(50 + r)).bfloat16()
                 for r in range(world)]
        full = comm.all_gather_lastdim(slabs[rank].to(dev))
        ok &= torch.equal(full.cpu(), oracle.all_gather_lastdim(slabs))
        # ring of sends: rank r -> r+1
        msg = torch.full((4096,), float(rank), dtype=torch.bfloat16, device=dev)
        got = torch.empty_like(msg)
        for _ in range(5):
            if rank % 2 == 0:
                comm.send(msg, (rank + 1) % world)
                comm.recv(got, (rank - 1) % world)
            else:
                comm.recv(got, (rank - 1) % world)
                comm.send(msg, (rank + 1) % world)
            torch.cuda.synchronize()
            ok &= bool((got.float() == float((rank - 1) % world)).all())
        comm.check_health()
        if rank == 0:
            print(f"[world {world}] NVLS multicast {'attached' if comm.nvls else 'NOT available: ' + getattr(comm, 'nvls_reason', '')}")
            for nline in notes:
                print(nline)
    finally:
        q.put((rank, bool(ok)))
        dist.barrier()
        comm.close()
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_real_multi_gpu(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok in res), res
