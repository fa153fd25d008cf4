"""The model-agnostic fusion pass on a hand-built post-grad FX graph (CPU): the pair
vllm.all_reduce -> vllm_ir.fused_add_rms_norm becomes one auto_functionalized call of
torch.ops.vllm.b200c_allreduce_add_rmsnorm; pairs it must not touch are left alone.  Mirrors what
AllReduceFusedAddRMSNormPattern matches (vllm/compilation/passes/fusion/allreduce_rms_fusion.py:347-416)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CODE = r'''
import sys, operator; sys.path.insert(0, %r)
import torch
from torch import fx
import vllm.ir.ops                                   # registers torch.ops.vllm_ir.*
import vllm.distributed.parallel_state               # registers torch.ops.vllm.all_reduce
from vllm_distributed_b200 import plugin, fusion_pass
plugin._register_fused_op()
from torch._higher_order_ops.auto_functionalize import auto_functionalized

AR = torch.ops.vllm.all_reduce.default
NORM = torch.ops.vllm_ir.fused_add_rms_norm.default

def build(variant):
    g = fx.Graph()
    x, res, w = g.placeholder("x"), g.placeholder("residual"), g.placeholder("weight")
    ar = g.call_function(AR, (x, "tp:0"))
    if variant == "two_users":
        other = g.call_function(torch.ops.aten.relu.default, (ar,))
    if variant == "var_size":
        n = g.call_function(NORM, (ar, res, w, 1e-5, 64))
    elif variant == "kwargs":
        n = g.call_function(NORM, (ar, res), dict(weight=w, epsilon=1e-6))
    else:
        n = g.call_function(NORM, (ar, res, w, 1e-5))
    y, r2 = g.call_function(operator.getitem, (n, 0)), g.call_function(operator.getitem, (n, 1))
    outs = (y, r2) if variant != "only_y" else (y,)
    if variant == "two_users":
        outs = outs + (other,)
    g.output(outs)
    return g

def targets(g):
    return [n.target for n in g.nodes if n.op == "call_function"]

for variant in ("plain", "kwargs", "only_y"):
    g = build(variant)
    assert fusion_pass.fuse_allreduce_rmsnorm(g) == 1, variant
    g.lint()
    t = targets(g)
    assert AR not in t and NORM not in t and auto_functionalized in t, (variant, t)
    fused = [n for n in g.nodes if n.target is auto_functionalized][0]
    assert fused.args == (torch.ops.vllm.b200c_allreduce_add_rmsnorm.default,)
    names = {k: (v.name if isinstance(v, fx.Node) else v) for k, v in fused.kwargs.items()}
    assert names["allreduce_in"] == "x" and names["residual"] == "residual" and names["weight"] == "weight"
    assert names["group_name"] == "tp:0" and names["eps"] == (1e-6 if variant == "kwargs" else 1e-5)
    out = [n for n in g.nodes if n.op == "output"][0].args[0]
    assert out[0].target is operator.getitem and out[0].args == (fused, 1)        # normalised output
    if variant != "only_y":
        assert out[1].args == (fused, 2)                                          # residual stream

for variant in ("two_users", "var_size"):
    g = build(variant)
    before = targets(g)
    assert fusion_pass.fuse_allreduce_rmsnorm(g) == 0 and targets(g) == before, variant

# the hook vLLM's config resolves: an attribute of the top-level package, callable on a graph
import vllm_distributed_b200
assert callable(vllm_distributed_b200.b200c_fusion_pass)
from vllm.config import CompilationConfig
cc = CompilationConfig(inductor_passes={"post_grad_custom_post_pass": "vllm_distributed_b200.b200c_fusion_pass"})
print("fusion pass ok")
'''


def test_fusion_pass_rewrites_the_pair_and_nothing_else():
    out = subprocess.run([sys.executable, "-c", CODE % ROOT], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-3000:]
    assert "fusion pass ok" in out.stdout
