"""GPU tests of the seams around the kernels: the vLLM IR-op provider (TP=1), the fused custom op
at TP=1, and what happens when a peer never shows up (bounded wait -> sticky error word, or a
trap that kills the context).  One GPU."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import oracle  # noqa: E402  (checker only)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bits_equal(a, b):
    return torch.equal(a.cpu().contiguous().view(torch.uint8), b.cpu().contiguous().view(torch.uint8))


@pytest.mark.parametrize("shape", [(64, 4096), (3, 896), (1300, 8192), (2500, 4096)])
def test_ir_op_dispatch_lands_in_libb200c(shape):
    """vllm.ir.ops.fused_add_rms_norm / rms_norm with provider "b200c" first: the dispatcher picks
    it for CUDA tensors and the result is this library's (residual bit-exact vs the oracle)."""
    assert torch.cuda.is_available(), "-m gpu tests need a CUDA device"
    from vllm import ir
    from vllm_distributed_b200 import norm_provider as npv
    assert npv.register(supported=True)
    g = torch.Generator().manual_seed(5)
    x = torch.randn(*shape, generator=g).bfloat16()
    r = torch.randn(*shape, generator=g).bfloat16()
    w = torch.randn(shape[1], generator=g).bfloat16()
    y_ref, r_ref = oracle.fused_add_rms_norm(x, r, w, 1e-5)
    xd, rd, wd = x.cuda(), r.cuda(), w.cuda()
    with ir.ops.fused_add_rms_norm.set_priority(["b200c", "native"]), \
            ir.ops.rms_norm.set_priority(["b200c", "native"]):
        assert ir.ops.fused_add_rms_norm.dispatch(xd, rd, wd, 1e-5).provider == "b200c"
        assert ir.ops.rms_norm.dispatch(xd, wd, 1e-5).provider == "b200c"
        # a strided view is not taken: the next provider serves it
        assert ir.ops.rms_norm.dispatch(xd[:, ::2], wd[::2], 1e-5).provider == "native"
        y, r2 = ir.ops.fused_add_rms_norm.maybe_inplace(xd, rd, wd, 1e-5)   # in place, as vLLM's layers call it
        assert y.data_ptr() == xd.data_ptr() and r2.data_ptr() == rd.data_ptr()
        y0 = ir.ops.rms_norm(x.cuda(), wd, 1e-6)
        # the functional overload must leave its inputs alone (inplace impls get clones)
        x_keep = x.cuda()
        y_f, r_f = ir.ops.fused_add_rms_norm(x_keep, r.cuda(), wd, 1e-5)
        assert _bits_equal(x_keep, x)
    assert _bits_equal(r2, r_ref) and _bits_equal(r_f, r_ref)
    torch.testing.assert_close(y.cpu().float(), y_ref.float(), atol=1e-3, rtol=1.6e-2)
    torch.testing.assert_close(y_f.cpu().float(), y_ref.float(), atol=1e-3, rtol=1.6e-2)
    y0_ref, _ = oracle.fused_add_rms_norm(x, None, w, 1e-6)
    torch.testing.assert_close(y0.cpu().float(), y0_ref.float(), atol=1e-3, rtol=1.6e-2)


def test_fused_custom_op_at_tp1_runs_the_local_norm():
    """torch.ops.vllm.b200c_allreduce_add_rmsnorm on a one-rank TP group (no communicator):
    the all-reduce is the identity, the op must still update both tensors in place."""
    code = r'''
import sys; sys.path.insert(0, %r)
import torch
from oracle import oracle
from vllm.config import VllmConfig, set_current_vllm_config
from vllm.distributed import parallel_state as ps
from vllm_distributed_b200 import plugin
plugin.register()
with set_current_vllm_config(VllmConfig()):
    ps.init_distributed_environment(world_size=1, rank=0, local_rank=0,
                                    distributed_init_method="tcp://127.0.0.1:29577", backend="nccl")
    ps.initialize_model_parallel(1, 1)
    tp = ps.get_tp_group()
    g = torch.Generator().manual_seed(3)
    x, r, w = (torch.randn(16, 4096, generator=g).bfloat16(), torch.randn(16, 4096, generator=g).bfloat16(),
               torch.randn(4096, generator=g).bfloat16())
    y_ref, r_ref = oracle.fused_add_rms_norm(x, r, w, 1e-5)
    xd, rd = x.cuda(), r.cuda()
    torch.ops.vllm.b200c_allreduce_add_rmsnorm(xd, rd, w.cuda(), 1e-5, tp.unique_name)
    torch.cuda.synchronize()
    assert torch.equal(rd.cpu(), r_ref), "residual not bit-exact"
    torch.testing.assert_close(xd.cpu().float(), y_ref.float(), atol=1e-3, rtol=1.6e-2)
print("tp1 op ok")
''' % ROOT
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, (out.stdout[-1500:], out.stderr[-3000:])
    assert "tp1 op ok" in out.stdout


def test_absent_peer_sets_the_sticky_error_word():
    """Rank 0 of a 2-rank group runs a REAL (not emulated) all-reduce while rank 1 never launches:
    the wait gives up after the configured timeout and b200c_query_error reports B200C_ETIMEOUT
    (the reference's barrier would spin forever, vllm_custom_all_reduce.cuh:190-211)."""
    from vllm_distributed_b200 import _cabi as C
    from vllm_distributed_b200.emu import EmuGroup
    g = EmuGroup(2, "cuda:0", max_bytes=1 << 20, ag_bytes=1 << 20, p2p_bytes=1 << 20)
    g.set_tuning(C.TUNE_SPIN_TIMEOUT_MS, 50)
    g.check_health()
    x = torch.randn(8, 1024, device="cuda").bfloat16()
    for variant in (C.ONESHOT, C.TWOSHOT):
        g.ranks[0].all_reduce(x, variant)      # the peer's pushes never come
    torch.cuda.synchronize()
    with pytest.raises(C.B200CError) as e:
        g.ranks[0].check_health()
    assert e.value.code == C.ETIMEOUT and "timed out" in str(e.value)
    g.ranks[1].check_health()                  # the other rank's state is its own
    g.close()


def test_absent_peer_with_trap_kills_the_context():
    """Same situation with B200C_TUNE_TRAP_ON_TIMEOUT (what B200Communicator sets): the kernel
    traps, the next CUDA call of the process fails -- a crashed worker, which the reference
    launcher turns into its fatal-disconnect path (launch.py:130-144) -- never wrong tokens."""
    code = r'''
import sys; sys.path.insert(0, %r)
import torch
from vllm_distributed_b200 import _cabi as C
from vllm_distributed_b200.emu import EmuGroup
g = EmuGroup(2, "cuda:0", max_bytes=1 << 20, ag_bytes=1 << 20, p2p_bytes=1 << 20)
g.set_tuning(C.TUNE_SPIN_TIMEOUT_MS, 50)
g.set_tuning(C.TUNE_TRAP_ON_TIMEOUT, 1)
x = torch.randn(8, 1024, device="cuda").bfloat16()
g.ranks[0].all_reduce(x, C.ONESHOT)
try:
    torch.cuda.synchronize()
    print("NO FAILURE")
except Exception as e:
    print("CUDA FAILED AS EXPECTED:", type(e).__name__)
import os; os._exit(0)
''' % ROOT
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert "CUDA FAILED AS EXPECTED" in out.stdout, (out.stdout[-1500:], out.stderr[-1500:])
