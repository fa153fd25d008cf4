"""model_patch._configure_model on fake layers (CPU): which row-parallel all-reduces are deferred
into the fused op, per pipeline stage -- the first layer of a stage receives a reduced tensor, the
last one hands a reduced tensor on, an un-fusable layer forces its predecessor to reduce."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CODE = r'''
import sys, types; sys.path.insert(0, %r)
import torch
from vllm_distributed_b200 import model_patch


class Lin:
    def __init__(self, bias=None): self.reduce_results, self.bias = True, bias
class Norm:
    def __init__(self, ok=True):
        self.weight = torch.ones(8); self.variance_epsilon = 1e-5
        self.has_weight = True; self.variance_size_override = None if ok else 4
class Layer:
    forward = model_patch._layer_forward      # a class whose forward IS the patched one
    def __init__(self, ok=True, bias=None):
        self.self_attn = types.SimpleNamespace(o_proj=Lin(bias))
        self.mlp = types.SimpleNamespace(down_proj=Lin())
        self.input_layernorm, self.post_attention_layernorm = Norm(), Norm(ok)
class Model:
    def __init__(self, layers, start, end): self.layers, self.start_layer, self.end_layer = layers, start, end

comm = types.SimpleNamespace(b200=object())
tp = types.SimpleNamespace(world_size=4, device_communicator=comm, unique_name="tp:0")
model_patch._tp = lambda: tp

# a 6-layer model, this stage owns layers 2..5
layers = [Layer() for _ in range(6)]
model_patch._configure_model(Model(layers, 2, 6))
mine = layers[2:6]
assert not any(getattr(l, "_b200c_fused", False) for l in layers[:2])      # other stage: untouched
assert all(l._b200c_fused and l._b200c_group == "tp:0" for l in mine)
assert [l._b200c_input_partial for l in mine] == [False, True, True, True]  # stage input is reduced
assert all(not l.self_attn.o_proj.reduce_results for l in mine)             # o_proj always deferred
assert [l.mlp.down_proj.reduce_results for l in mine] == [False, False, False, True]  # last reduces

# an un-fusable layer in the middle (norm with a variance override): its predecessor must reduce
layers = [Layer(), Layer(), Layer(ok=False), Layer()]
model_patch._configure_model(Model(layers, 0, 4))
assert [getattr(l, "_b200c_fused", False) for l in layers] == [True, True, False, True]
assert layers[1].mlp.down_proj.reduce_results          # feeds the un-fused layer: reduced
assert layers[2].self_attn.o_proj.reduce_results and layers[2].mlp.down_proj.reduce_results
assert layers[3]._b200c_input_partial is False          # receives layer 2's reduced output
assert layers[0].mlp.down_proj.reduce_results is False and layers[1]._b200c_input_partial is True

# a projection with a bias is never deferred (rank 0 alone adds it: linear.py:1551-1553)
layers = [Layer(bias=torch.zeros(8)), Layer()]
model_patch._configure_model(Model(layers, 0, 2))
assert not getattr(layers[0], "_b200c_fused", False) and layers[1]._b200c_fused
assert layers[1]._b200c_input_partial is False

# a subclass layer with its OWN forward (Qwen3 / Mistral / Glm4 style) is never touched, and its
# fused neighbours reduce around it
class OwnForward(Layer):
    def forward(self, *a, **k): raise AssertionError("not called")
layers = [Layer(), OwnForward(), Layer(), Layer()]
model_patch._configure_model(Model(layers, 0, 4))
assert [getattr(l, "_b200c_fused", False) for l in layers] == [True, False, True, True]
assert layers[1].self_attn.o_proj.reduce_results and layers[1].mlp.down_proj.reduce_results
assert layers[0].mlp.down_proj.reduce_results and not layers[2]._b200c_input_partial
# ... while a subclass that inherits the patched forward is fused like its parent
class Inherits(Layer): pass
layers = [Inherits(), Inherits()]
model_patch._configure_model(Model(layers, 0, 2))
assert all(l._b200c_fused for l in layers) and layers[1]._b200c_input_partial

# EAGLE3 aux hidden states read hidden+residual after each layer: no fusion at all
layers = [Layer(), Layer()]
m = Model(layers, 0, 2); m.aux_hidden_state_layers = (1,)
model_patch._configure_model(m)
assert not any(getattr(l, "_b200c_fused", False) for l in layers)
assert all(l.self_attn.o_proj.reduce_results and l.mlp.down_proj.reduce_results for l in layers)

# all-or-nothing: a failure half way puts every flag back
class Exploding(Lin):
    def __setattr__(self, k, v):
        if k == "reduce_results" and v is False and getattr(self, "armed", False): raise RuntimeError("boom")
        object.__setattr__(self, k, v)
layers = [Layer(), Layer(), Layer()]
layers[2].self_attn.o_proj = Exploding(); layers[2].self_attn.o_proj.armed = True
try:
    model_patch._configure_model(Model(layers, 0, 3)); raise SystemExit("expected a failure")
except RuntimeError:
    pass
assert not any(getattr(l, "_b200c_fused", False) for l in layers)
assert all(l.self_attn.o_proj.reduce_results and l.mlp.down_proj.reduce_results for l in layers)

# TP=1 or a foreign communicator: nothing is touched
for bad in (types.SimpleNamespace(world_size=1, device_communicator=None, unique_name="tp:0"),
            types.SimpleNamespace(world_size=4, device_communicator=types.SimpleNamespace(), unique_name="tp:0")):
    model_patch._tp = lambda bad=bad: bad
    layers = [Layer(), Layer()]
    model_patch._configure_model(Model(layers, 0, 2))
    assert not any(getattr(l, "_b200c_fused", False) for l in layers)
    assert all(l.self_attn.o_proj.reduce_results and l.mlp.down_proj.reduce_results for l in layers)
print("model patch ok")
'''


def test_layer_fusion_plan():
    from _shared_vllm_run import run_all
    r = run_all()
    assert "model patch ok" in r["out"], r["err"][-3000:]
