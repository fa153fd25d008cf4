"""The TP=1 seam on CPU: libb200c's norm kernels register as provider "b200c" of vLLM's IR ops
(vllm/kernels/vllm_c.py:22-91 is the stock provider they sit in front of), the dispatcher skips
them for tensors they do not take (CPU tensors here) and falls through to the native definition,
and the platform's default priority list gets "b200c" in front of "vllm_c" only where vLLM itself
would have used its C kernels (vllm/platforms/cuda.py:574-593)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CODE = r'''
import sys, types; sys.path.insert(0, %r)
import torch
from vllm import ir
from vllm_distributed_b200 import norm_provider as npv

assert npv.register(supported=True) and npv.register(supported=True)          # idempotent
for op in (ir.ops.rms_norm, ir.ops.fused_add_rms_norm):
    assert "b200c" in op.impls and op.impls["b200c"].supported
assert ir.ops.fused_add_rms_norm.impls["b200c"].inplace and not ir.ops.rms_norm.impls["b200c"].inplace

# CPU tensors are not taken: dispatch falls through to native, results are the native ones
x, r, w = torch.randn(3, 64).bfloat16(), torch.randn(3, 64).bfloat16(), torch.randn(64).bfloat16()
with ir.ops.fused_add_rms_norm.set_priority(["b200c", "native"]), ir.ops.rms_norm.set_priority(["b200c", "native"]):
    assert ir.ops.fused_add_rms_norm.dispatch(x, r, w, 1e-5).provider == "native"
    assert ir.ops.rms_norm.dispatch(x, w, 1e-5).provider == "native"
    y, r2 = ir.ops.fused_add_rms_norm(x, r, w, 1e-5)
y_ref, r_ref = ir.ops.fused_add_rms_norm.impls["native"].impl_fn(x, r, w, 1e-5)
assert torch.equal(y, y_ref) and torch.equal(r2, r_ref)
# the argument filter itself (what a CUDA tensor would have to satisfy)
take = ir.ops.fused_add_rms_norm.impls["b200c"].supports_args
assert not take(x, r, w, 1e-5) and not take(x, r, w, 1e-5, 32)

# priority: in front of vllm_c when vLLM would use its C kernels, untouched under Inductor
class Prio:
    def __init__(self, lst): self.rms_norm, self.fused_add_rms_norm = list(lst), list(lst)
class Platform:
    mode = ["vllm_c", "native"]
    @classmethod
    def get_default_ir_op_priority(cls, vllm_config): return Prio(cls.mode)
npv.prioritise(Platform); npv.prioritise(Platform)                              # idempotent
p = Platform.get_default_ir_op_priority(None)
assert p.rms_norm == ["b200c", "vllm_c", "native"] and p.fused_add_rms_norm == ["b200c", "vllm_c", "native"]
Platform.mode = ["native"]
p = Platform.get_default_ir_op_priority(None)
assert p.rms_norm == ["native"] and p.fused_add_rms_norm == ["native"]
Platform.mode = ["oink", "vllm_c", "native"]
assert Platform.get_default_ir_op_priority(None).rms_norm == ["oink", "b200c", "vllm_c", "native"]
print("norm provider ok")
'''


def test_norm_provider_registration_dispatch_and_priority():
    from _shared_vllm_run import run_all
    r = run_all()
    assert "norm provider ok" in r["out"], r["err"][-3000:]
