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
        # AUTO thresholds from a table measured on this box: every rank must choose the same
        # ones, and AUTO must stay bit-exact afterwards
        tab = comm.calibrate(cpu_group, hidden=4096, iters=4, sizes=(64 << 10, 1 << 20, 4 << 20))
        chosen = [None] * world
        dist.all_gather_object(chosen, tab["chosen"], group=cpu_group)
        ok &= all(c == chosen[0] for c in chosen)
        notes.append(f"calibrated: {tab['chosen']}")
        xs = [torch.randn(300, 4096, generator=torch.Generator().manual_seed(1234 + r)).bfloat16() for r in range(world)]
        ok &= torch.equal(comm.all_reduce(xs[rank].to(dev)).cpu(), oracle.all_reduce(xs))
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


def _vllm_worker(rank, world, port, q):
    """The vLLM-facing layer on real GPUs: plugin.register() -> GroupCoordinator builds a
    B200Communicator -> tensor_model_parallel_all_reduce / all_gather and the fused custom op."""
    ok, why = True, ""
    try:
        torch.cuda.set_device(rank)
        from oracle import oracle
        from vllm.config import VllmConfig, set_current_vllm_config
        from vllm.distributed import parallel_state as ps
        from vllm.distributed.communication_op import (tensor_model_parallel_all_gather,
                                                       tensor_model_parallel_all_reduce)
        from vllm_distributed_b200 import plugin
        from vllm_distributed_b200.communicator import B200Communicator
        plugin.register()
        with set_current_vllm_config(VllmConfig()):
            ps.init_distributed_environment(world_size=world, rank=rank, local_rank=rank,
                                            distributed_init_method=f"tcp://127.0.0.1:{port}",
                                            backend="nccl")
            ps.initialize_model_parallel(world, 1)
            tp = ps.get_tp_group()
            comm = tp.device_communicator
            assert isinstance(comm, B200Communicator) and comm.b200 is not None, type(comm)
            for t, h in [(1, 4096), (64, 8192), (300, 4096)]:
                xs = [torch.randn(t, h, generator=torch.Generator().manual_seed(1234 + r)).bfloat16()
                      for r in range(world)]
                res = torch.randn(t, h, generator=torch.Generator().manual_seed(7)).bfloat16()
                gamma = torch.randn(h, generator=torch.Generator().manual_seed(8)).bfloat16()
                y_ref, r_ref, ar_ref = oracle.all_reduce_add_rms_norm(xs, res, gamma, 1e-5)
                ar = tensor_model_parallel_all_reduce(xs[rank].cuda())
                xd, rd = xs[rank].cuda(), res.cuda()
                torch.ops.vllm.b200c_allreduce_add_rmsnorm(xd, rd, gamma.cuda(), 1e-5, tp.unique_name)
                torch.cuda.synchronize()
                ok &= torch.equal(ar.cpu(), ar_ref) and torch.equal(rd.cpu(), r_ref)
                ok &= bool(torch.allclose(xd.cpu().float(), y_ref.float(), atol=1e-3, rtol=1.6e-2))
            slabs = [torch.randn(5, 1024, generator=torch.Generator().manual_seed(50 + r)).bfloat16()
                     for r in range(world)]
            full = tensor_model_parallel_all_gather(slabs[rank].cuda(), -1)
            ok &= torch.equal(full.cpu(), oracle.all_gather_lastdim(slabs))
            comm.check_health()
            torch.distributed.barrier()
        ps.destroy_model_parallel()
        ps.destroy_distributed_environment()
    except Exception:  # noqa: BLE001
        import traceback
        ok, why = False, traceback.format_exc()[-1500:]
    q.put((rank, bool(ok), why))


@pytest.mark.parametrize("world", [2, 8])
def test_vllm_communicator_api_on_real_gpus(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_vllm_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), res
