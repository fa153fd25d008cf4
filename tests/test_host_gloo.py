"""Host-side logic of the N>1 path on CPU: world_size 2 and 4 over gloo (no GPU, no compute calls
into libb200c).  Covers the IPC-handle exchange, the PP tensor-dict channel (against the golden
fixtures written by vLLM's own send/recv_tensor_dict) and bench.py's reference arm."""
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(fn, world, *args):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=fn, args=(r, world, port, q) + args) for r in range(world)]
    for p in procs:
        p.start()
    out = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    return dict(out)


# ---- arena handle exchange ---------------------------------------------------------------------

def _exchange_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from vllm_distributed_b200 import comm as bc
    dist.init_process_group("gloo", rank=rank, world_size=world,
                            init_method=f"tcp://127.0.0.1:{port}")
    seen = {}
    bc.arena_bytes = lambda w, a, b, c: 4096
    bc.alloc_shared = lambda n: (1000 + rank, bytes([rank]) * 64)
    bc.open_handle = lambda h: 2000 + h[0]

    def fake_init(self, r, w, ptrs, max_bytes, ag_bytes, p2p_bytes, device, owned_arena=0, opened=None):
        seen.update(rank=r, world=w, ptrs=list(ptrs), owned=owned_arena, opened=list(opened or []))
        self.ctx = None
        self._owned_arena = 0
        self._opened = []
    bc.B200Comm.__init__ = fake_init
    bc.B200Comm.close = lambda self: None

    class _NoDev:
        def __enter__(self): return self
        def __exit__(self, *a): return False
    torch.cuda.device = lambda d: _NoDev()
    bc.B200Comm.from_process_group(dist.group.WORLD, torch.device("cpu"), 1 << 20, 1 << 20, 1 << 20)
    q.put((rank, seen))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_ipc_handle_exchange_orders_arenas_by_rank(world):
    res = _spawn(_exchange_worker, world)
    for rank, seen in res.items():
        assert seen["rank"] == rank and seen["world"] == world
        want = [1000 + r if r == rank else 2000 + r for r in range(world)]
        assert seen["ptrs"] == want          # own arena at [rank], peers' mappings elsewhere
        assert seen["owned"] == 1000 + rank
        assert sorted(seen["opened"]) == sorted(p for p in want if p >= 2000)


# ---- PP channel ------------------------------------------------------------------------------------

class _FakeComm:
    """Stands in for B200Communicator on a CPU box: CUDA tensors would go to send/recv."""
    b200 = object()

    def send(self, t, dst):
        raise AssertionError("CPU tensors must not reach the device mailbox")

    def recv(self, size, dtype, src):
        raise AssertionError("CPU tensors must not reach the device mailbox")


def _pp_worker(rank, world, port, q, tp, pp, golden):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    from oracle import oracle
    from vllm.config import DeviceConfig, VllmConfig, set_current_vllm_config
    from vllm.distributed import parallel_state as ps
    from vllm_distributed_b200 import pp_channel
    g = dict(np.load(golden))
    got = {}
    with set_current_vllm_config(VllmConfig(device_config=DeviceConfig(device="cpu"))):
        ps.init_distributed_environment(world_size=world, rank=rank, local_rank=rank,
                                        distributed_init_method=f"tcp://127.0.0.1:{port}",
                                        backend="gloo")
        ps.initialize_model_parallel(tp, pp)
        pp_channel.apply()
        ppg = ps.get_pp_group()
        ppg.device_communicator = _FakeComm()   # route through the patched channel
        tpg = ps.get_tp_group() if tp > 1 else None
        hs, rs = oracle.from_bits(g["hidden_states_in"]), oracle.from_bits(g["residual_in"])
        paths = []
        orig_send_meta = pp_channel.send_meta
        pp_channel.send_meta = lambda grp, meta, dst: paths.append(orig_send_meta(grp, meta, dst))
        # several steps: the channel must stay in lock-step.  Steps 0-2 change the metadata
        # every time, steps 3-5 repeat it (the decode steady state), step 6 changes it again,
        # and 20 more repeats wrap the 16-slot ring.
        tags = [0, 1, 2, 99, 99, 99, 7] + [7] * 20
        for step, tag in enumerate(tags):
            if ppg.rank_in_group == 0:
                ppg.send_tensor_dict({"hidden_states": hs, "residual": rs, "step": tag},
                                     all_gather_group=tpg)
            else:
                d = ppg.recv_tensor_dict(all_gather_group=tpg)
                assert d["step"] == tag
                got = {k: oracle.to_bits(v) for k, v in d.items() if isinstance(v, torch.Tensor)}
        if ppg.rank_in_group == 0:
            # only the first message of the pair pays the reference's pickled gloo handshake
            assert paths == ["gloo", "shm", "shm", "shm", "shm-same", "shm-same", "shm"] + ["shm-same"] * 20, paths
        torch.distributed.barrier()
    q.put((rank, (ppg.rank_in_group, ps.get_tp_group().rank_in_group, got)))
    ps.destroy_model_parallel()
    ps.destroy_distributed_environment()


@pytest.mark.parametrize("name,tp,pp", [("pp2_tp1", 1, 2), ("pp2_tp2", 2, 2)])
def test_pp_channel_matches_vllm_golden(golden_dir, name, tp, pp):
    golden = os.path.join(golden_dir, name + ".npz")
    g = dict(np.load(golden))
    res = _spawn(_pp_worker, tp * pp, tp, pp, golden)
    receivers = 0
    for rank, (pp_rank, tp_rank, got) in res.items():
        if pp_rank == 1:
            receivers += 1
            assert np.array_equal(got["hidden_states"], g[f"hidden_states_out_tp{tp_rank}"])
            assert np.array_equal(got["residual"], g[f"residual_out_tp{tp_rank}"])
    assert receivers == tp


def _nvls_attach_worker(rank, world, port, q):
    """B200Comm.try_attach_nvls with torch's symmetric memory and the C library mocked: the
    group attaches the in-switch buffer only if EVERY rank got a multicast pointer."""
    sys.path.insert(0, ROOT)
    import types
    import torch.distributed as dist
    import torch.distributed._symmetric_memory as symm_mem
    from vllm_distributed_b200 import _cabi as C
    from vllm_distributed_b200 import comm as bc
    dist.init_process_group("gloo", rank=rank, world_size=world, init_method=f"tcp://127.0.0.1:{port}")
    calls = []

    class FakeLib:
        def b200c_nvls_bytes(self, max_bytes, out):
            out._obj.value = 32 * 1024 + max_bytes + 256
            return 0

        def b200c_attach_nvls(self, ctx, uc, mc, nbytes):
            calls.append((uc, mc, nbytes))
            return 0

        def b200c_set_tuning(self, ctx, key, value):
            calls.append(("tune", key, value))
            return 0
    C._lib = FakeLib()

    class _NoDev:
        def __enter__(self): return self
        def __exit__(self, *a): return False
    torch.cuda.device = lambda d: _NoDev()
    outcomes = []
    for mc_of_rank in ({0: 0x7000, 1: 0}, {0: 0x7000, 1: 0x7100}):
        symm_mem.enable_symm_mem_for_group = lambda name: None
        symm_mem.empty = lambda n, dtype=None, device=None: types.SimpleNamespace(data_ptr=lambda: 0x1000 + rank, numel=n)
        symm_mem.rendezvous = lambda buf, name, mc=mc_of_rank: types.SimpleNamespace(multicast_ptr=mc[rank])
        c = object.__new__(bc.B200Comm)
        c.rank, c.world, c.device, c.max_bytes, c.ctx = rank, world, torch.device("cpu"), 1 << 20, None
        c.nvls, c._nvls_buf = False, None
        calls.clear()
        ok = c.try_attach_nvls(dist.group.WORLD, types.SimpleNamespace(group_name="g0"))
        outcomes.append((ok, c.nvls, list(calls)))
    bc.B200Comm.__del__ = lambda self: None
    q.put((rank, outcomes))
    dist.barrier()
    dist.destroy_process_group()


def test_nvls_attach_is_all_or_nothing():
    res = _spawn(_nvls_attach_worker, 2)
    for rank, ((ok1, nvls1, calls1), (ok2, nvls2, calls2)) in res.items():
        assert (ok1, nvls1, calls1) == (False, False, [])            # rank 1 had no multicast pointer: nobody attaches
        assert ok2 and nvls2 and calls2 == [(0x1000 + rank, 0x7000 + 0x100 * rank, 32 * 1024 + (1 << 20) + 256)]


def _ring_worker(rank, world, port, q):
    """pp_channel's shared-memory object ring on its own (no vLLM): first message over the
    group's send_object, then the ring -- wrap-around with a slow reader, a payload too large for
    a slot (falls back to send_object for that message), and the same-as-before marker."""
    sys.path.insert(0, ROOT)
    import pickle
    import time
    import torch.distributed as dist
    from vllm_distributed_b200 import pp_channel
    dist.init_process_group("gloo", rank=rank, world_size=world, init_method=f"tcp://127.0.0.1:{port}")

    class Group:
        rank_in_group = rank

        def send_object(self, obj, dst):
            t = torch.frombuffer(bytearray(pickle.dumps(obj)), dtype=torch.uint8)
            dist.send(torch.tensor([t.numel()]), dst=dst)
            dist.send(t, dst=dst)

        def recv_object(self, src):
            n = torch.empty(1, dtype=torch.long)
            dist.recv(n, src=src)
            t = torch.empty(int(n.item()), dtype=torch.uint8)
            dist.recv(t, src=src)
            return pickle.loads(t.numpy().tobytes())

    g = Group()
    big = "x" * (pp_channel.ShmObjectChannel.SLOT_BYTES + 100)
    msgs = [("first", 0)] + [("m", i) for i in range(40)] + [("big", big), ("after", 1), ("after", 1), ("after", 1)]
    paths, got = [], []
    if rank == 0:
        for m in msgs:
            paths.append(pp_channel.send_meta(g, m, 1))
        shm_file = g._b200c_meta_channels[("send", 1)].path
    else:
        for i, _ in enumerate(msgs):
            if i == 5:
                time.sleep(0.5)          # the writer runs 16 slots ahead and has to wait for acks
            got.append(pp_channel.recv_meta(g, 0))
        shm_file = None
    dist.barrier()
    q.put((rank, (paths, got == msgs, shm_file is None or not os.path.exists(shm_file))))
    dist.destroy_process_group()


def test_pp_metadata_ring_wraps_falls_back_and_repeats():
    res = _spawn(_ring_worker, 2)
    paths, _, unlinked = res[0]
    assert paths[0] == "gloo" and set(paths[1:41]) == {"shm"} and len(paths) == 45
    assert paths[41] == "gloo"                      # too large for a slot: that one message over gloo
    assert paths[42:] == ["shm", "shm-same", "shm-same"]
    assert unlinked                                  # the /dev/shm file is gone once both sides mapped it
    assert res[1][1]                                 # the receiver saw every message, in order


# ---- bench.py reference arm ------------------------------------------------------------------------

def test_bench_reference_arm_prints_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference",
                          "--gpus", "2", "--steps", "1", "--warmup", "1", "--batch", "2"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    r = json.loads(lines[0])
    assert r["impl"] == "reference" and r["n_gpus"] == 2 and r["value"] > 0
    assert r["cpu_baseline"]["kind"] in ("reference", "port") and r["cpu_baseline"]["cores"] >= 1
    assert r["e2e"]["h2d_bytes_per_step"] == 0 and r["e2e"]["d2h_bytes_per_step"] == 0
    for k in ("metric", "unit", "steps", "warmup", "ms_per_step", "higher_is_better", "config"):
        assert k in r
