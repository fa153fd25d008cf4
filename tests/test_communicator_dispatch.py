"""Dispatch rules of B200Communicator on CPU: what goes to the NVLink arenas, what falls back to
the stock communicator (the parent class), with the parent and the arena object mocked out.
Mirrors the fall-through behaviour of vLLM's CudaCommunicator.all_reduce
(vllm/distributed/device_communicators/cuda_communicator.py:241-298)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CODE = r'''
import sys; sys.path.insert(0, %r)
import torch
from vllm.distributed.device_communicators.cuda_communicator import CudaCommunicator
from vllm_distributed_b200.communicator import B200Communicator

calls = []
CudaCommunicator.all_reduce = lambda self, t: calls.append(("super.all_reduce", tuple(t.shape))) or t
CudaCommunicator.all_gather = lambda self, t, dim=-1: calls.append(("super.all_gather", dim)) or t
CudaCommunicator.send = lambda self, t, dst=None: calls.append(("super.send", dst))
CudaCommunicator.recv = lambda self, size, dtype, src=None: calls.append(("super.recv", src)) or torch.empty(size, dtype=dtype)
CudaCommunicator.destroy = lambda self: calls.append(("super.destroy",))


class FakeArena:
    world = 4
    def __init__(self): self.closed = False
    def can_all_reduce(self, x): return x.is_contiguous() and (x.numel() * x.element_size()) %% 16 == 0
    def can_fuse_norm(self, x): return self.can_all_reduce(x)
    def can_all_gather_lastdim(self, x): return x.is_contiguous() and (x.shape[-1] * x.element_size()) %% 4 == 0
    def can_p2p(self, t): return t.is_contiguous() and (t.numel() * t.element_size()) %% 16 == 0
    def all_reduce(self, x): calls.append(("arena.all_reduce", tuple(x.shape))); return x.clone()
    def all_reduce_add_rmsnorm(self, x, r, w, eps, inplace=False):
        calls.append(("arena.fused", inplace)); return x, r
    def all_gather_lastdim(self, x):
        calls.append(("arena.all_gather_lastdim", tuple(x.shape)))
        return torch.cat([x] * self.world, dim=-1)
    def send(self, t, dst): calls.append(("arena.send", dst))
    def recv(self, t, src): calls.append(("arena.recv", src)); return t
    def check_health(self): calls.append(("arena.check_health",))
    def close(self): self.closed = True


c = object.__new__(B200Communicator)
c.b200 = FakeArena(); c.world_size = 4; c.rank_in_group = 1; c.device = torch.device("cpu")

# all-reduce: eligible tensors go to the arena, odd sizes / strided views fall through
c.all_reduce(torch.zeros(4, 64, dtype=torch.bfloat16))
c.all_reduce(torch.zeros(3, 5, dtype=torch.bfloat16))            # 30 bytes: not a multiple of 16
c.all_reduce(torch.zeros(8, 64, dtype=torch.bfloat16)[:, ::2])    # not contiguous
assert [k for k, *_ in calls] == ["arena.all_reduce", "super.all_reduce", "super.all_reduce"], calls

# all-gather: last dim (also when given as a positive index) and the dim-0 PP re-assembly
calls.clear()
out = c.all_gather(torch.zeros(2, 8), dim=-1); assert out.shape == (2, 32)
out = c.all_gather(torch.zeros(2, 8), dim=1); assert out.shape == (2, 32)
out = c.all_gather(torch.zeros(6, 8), dim=0); assert out.shape == (24, 8)
c.all_gather(torch.zeros(2, 3, 8), dim=1)                         # middle dim: stock path
assert [k for k, *_ in calls] == ["arena.all_gather_lastdim"] * 3 + ["super.all_gather"], calls
assert calls[2][1] == (1, 48)                                     # dim 0 == one flattened row

# send / recv: defaults are the ring neighbours, like the reference communicator
calls.clear()
c.send(torch.zeros(64, dtype=torch.bfloat16)); c.send(torch.zeros(3, dtype=torch.bfloat16), 3)
t = c.recv(torch.Size([64]), torch.bfloat16); c.recv(torch.Size([3]), torch.bfloat16, 0)
assert calls == [("arena.send", 2), ("super.send", 3), ("arena.recv", 0), ("super.recv", 0)], calls

# fused op: arena when eligible; otherwise stock all-reduce + this library's norm (not reached on CPU)
calls.clear()
x = torch.zeros(4, 64, dtype=torch.bfloat16); r = torch.zeros_like(x); w = torch.ones(64, dtype=torch.bfloat16)
c.all_reduce_add_rmsnorm(x, r, w, 1e-5, inplace=True)
assert calls == [("arena.fused", True)], calls

# no arena (multi-node group, disabled, ...): everything is the stock communicator
calls.clear(); arena = c.b200; c.b200 = None
c.all_reduce(torch.zeros(4, 64, dtype=torch.bfloat16)); c.all_gather(torch.zeros(2, 8)); c.send(torch.zeros(64))
assert [k for k, *_ in calls] == ["super.all_reduce", "super.all_gather", "super.send"], calls

# destroy releases the arenas after surfacing device-side errors, then the parent's resources
calls.clear(); c.b200 = arena
c.destroy()
assert arena.closed and c.b200 is None and calls == [("arena.check_health",), ("super.destroy",)], calls
print("dispatch ok")
'''


def test_variant_thresholds_from_a_measured_table():
    """choose_thresholds: the AUTO crossovers as a pure function of a measured table."""
    from vllm_distributed_b200.comm import choose_thresholds
    K, M = 1 << 10, 1 << 20
    # the TP=8 sweep of profiles/r2_sweep_tp8_final.json, abridged
    table = {16 * K: {"oneshot": 6.25, "twoshot": 9.18}, 256 * K: {"oneshot": 8.77, "twoshot": 11.08},
             1 * M: {"oneshot": 17.91, "twoshot": 12.44}, 4 * M: {"oneshot": 61.9, "twoshot": 21.8, "nvls": 33.6},
             32 * M: {"twoshot": 119.3, "nvls": 134.8}, 64 * M: {"twoshot": 218.8, "nvls": 197.6}}
    assert choose_thresholds(table) == {"oneshot_max_bytes": 256 * K, "nvls_min_bytes": 64 * M}
    # one-shot never wins / NVLS never wins / a noisy one-shot win above the crossover is ignored
    assert choose_thresholds({1 * M: {"oneshot": 9.0, "twoshot": 8.0}}) == {"oneshot_max_bytes": 0, "nvls_min_bytes": -1}
    noisy = {64 * K: {"oneshot": 5.0, "twoshot": 7.0}, 1 * M: {"oneshot": 9.0, "twoshot": 8.0},
             2 * M: {"oneshot": 8.9, "twoshot": 9.0}}
    assert choose_thresholds(noisy)["oneshot_max_bytes"] == 64 * K
    # NVLS must win at EVERY larger size to be chosen from a size on
    gap = {8 * M: {"twoshot": 40.0, "nvls": 39.0}, 32 * M: {"twoshot": 100.0, "nvls": 110.0},
           64 * M: {"twoshot": 200.0, "nvls": 180.0}}
    assert choose_thresholds(gap)["nvls_min_bytes"] == 64 * M


def test_dispatch_rules():
    from _shared_vllm_run import run_all
    r = run_all()
    assert "dispatch ok" in r["out"], r["err"][-3000:]
