"""Eager/compiled fused path for the decoder families BASELINE.json names (Llama, Qwen2).

The reference gets its fused all-reduce + residual + RMSNorm from a torch.compile pattern
rewrite (vllm/compilation/passes/fusion/allreduce_rms_fusion.py:347-416).  Here the same pairing
is made at the layer: the row-parallel projections (o_proj, down_proj;
vllm/model_executor/layers/linear.py:1558-1561) keep their partial sums, and the RMSNorm that
follows (vllm/model_executor/models/llama.py:316-333, qwen2.py:292-312) is replaced by ONE call of
torch.ops.vllm.b200c_allreduce_add_rmsnorm, which works eagerly, under Dynamo and inside CUDA
graphs.  The last layer of every pipeline stage still reduces normally, so the final norm and
the PP hand-off (llama.py:425-430) see ordinary tensors.
"""
from __future__ import annotations

import functools

import torch

_APPLIED = False


def _tp():
    from vllm.distributed.parallel_state import get_tp_group
    return get_tp_group()


def _fusable_norm(norm) -> bool:
    return (getattr(norm, "has_weight", True) and getattr(norm, "variance_size_override", None) is None
            and isinstance(getattr(norm, "weight", None), torch.Tensor))


def _layer_forward(self, positions, hidden_states, residual, **kwargs):
    if not getattr(self, "_b200c_fused", False):
        return self._b200c_orig_forward(positions, hidden_states, residual, **kwargs)
    group_name = self._b200c_group
    if residual is None:
        residual = hidden_states
        hidden_states = self.input_layernorm(hidden_states)
    elif self._b200c_input_partial:
        w = self.input_layernorm
        torch.ops.vllm.b200c_allreduce_add_rmsnorm(hidden_states, residual, w.weight.data,
                                                   w.variance_epsilon, group_name)
    else:
        hidden_states, residual = self.input_layernorm(hidden_states, residual)
    hidden_states = self.self_attn(positions=positions, hidden_states=hidden_states, **kwargs)
    w = self.post_attention_layernorm
    torch.ops.vllm.b200c_allreduce_add_rmsnorm(hidden_states, residual, w.weight.data,
                                               w.variance_epsilon, group_name)
    hidden_states = self.mlp(hidden_states)
    return hidden_states, residual


def _plan(model, layers) -> list[dict] | None:
    """Decide, without touching anything, what each layer of this stage would become.
    A layer is fused only if it RUNS the patched forward: its class is one of the patched
    decoder-layer classes or a subclass that does not override forward.  Subclass models build
    other layer classes through the wrapped constructors -- Qwen3Model(Qwen2Model) with
    Qwen3DecoderLayer, MistralModel(LlamaModel) with its own forward, Glm4 (which normalises the
    o_proj output BEFORE the residual add, so deferring its all-reduce would be wrong even with a
    patched forward) -- and must keep reduce_results=True."""
    if getattr(model, "aux_hidden_state_layers", None):
        # EAGLE3 reads hidden_states + residual after each layer (llama.py _maybe_add_hidden_state):
        # a deferred all-reduce would hand it partial sums
        return None
    plan = []
    for i, layer in enumerate(layers):
        ok = (getattr(type(layer), "forward", None) is _layer_forward
              and hasattr(layer, "self_attn") and hasattr(layer.self_attn, "o_proj")
              and hasattr(layer, "mlp") and hasattr(layer.mlp, "down_proj")
              and _fusable_norm(layer.input_layernorm)
              and _fusable_norm(layer.post_attention_layernorm)
              and layer.self_attn.o_proj.reduce_results and layer.mlp.down_proj.reduce_results
              and layer.self_attn.o_proj.bias is None and layer.mlp.down_proj.bias is None)
        plan.append({"fused": ok})
    for i, step in enumerate(plan):
        if not step["fused"]:
            continue
        nxt_fused = i + 1 < len(plan) and plan[i + 1]["fused"]
        # the last layer of the stage hands a reduced tensor to the final norm / the PP send; a
        # layer in front of an un-fused one must reduce too
        step["down_reduce"] = not nxt_fused
        step["input_partial"] = i > 0 and plan[i - 1]["fused"] and not plan[i - 1]["down_reduce"]
    return plan


def _configure_model(model) -> None:
    """Decide, per layer of this pipeline stage, which all-reduces are deferred into the norm,
    then apply the plan all-or-nothing: on any exception every flag already flipped is put back,
    so a model is never left with a dropped all-reduce."""
    tp = _tp()
    comm = tp.device_communicator
    if tp.world_size < 2 or comm is None or getattr(comm, "b200", None) is None:
        return
    start, end = getattr(model, "start_layer", 0), getattr(model, "end_layer", len(model.layers))
    layers = list(model.layers)[start:end]
    plan = _plan(model, layers)
    if plan is None:
        return
    undo = []
    try:
        for layer, step in zip(layers, plan):
            if not step["fused"]:
                continue
            o, d = layer.self_attn.o_proj, layer.mlp.down_proj
            undo.append((layer, o.reduce_results, d.reduce_results))
            layer._b200c_group = tp.unique_name
            layer._b200c_input_partial = step["input_partial"]
            o.reduce_results = False
            d.reduce_results = step["down_reduce"]
            layer._b200c_fused = True  # last: the patched forward only trusts a complete layer
    except Exception:
        for layer, o_rr, d_rr in undo:
            layer._b200c_fused = False
            layer.self_attn.o_proj.reduce_results = o_rr
            layer.mlp.down_proj.reduce_results = d_rr
        raise
    from vllm.logger import init_logger
    init_logger("vllm.b200c.model_patch").info(
        "b200c: fused all-reduce+residual+RMSNorm on %d of %d decoder layers (group %s)",
        sum(1 for st in plan if st["fused"]), len(layers), tp.unique_name)


def _wrap_model_init(model_cls) -> None:
    if getattr(model_cls.__init__, "_b200c_wrapped", False):
        return
    orig_init = model_cls.__init__

    @functools.wraps(orig_init)
    def init(self, *args, **kwargs):
        orig_init(self, *args, **kwargs)
        if getattr(self, "_b200c_configured", False):
            return  # a subclass constructor chained through several wrapped __init__s
        self._b200c_configured = True
        try:
            _configure_model(self)
        except Exception:
            # _configure_model has rolled everything back: the model runs un-fused, correctly
            import logging
            logging.getLogger(__name__).exception("b200c: layer fusion disabled for %s", type(self).__name__)
    init._b200c_wrapped = True
    model_cls.__init__ = init


def apply() -> None:
    global _APPLIED
    if _APPLIED:
        return
    _APPLIED = True
    from vllm.model_executor.models import llama, qwen2
    for mod, layer_name, model_name in ((llama, "LlamaDecoderLayer", "LlamaModel"),
                                        (qwen2, "Qwen2DecoderLayer", "Qwen2Model")):
        layer_cls = getattr(mod, layer_name)
        if not hasattr(layer_cls, "_b200c_orig_forward"):
            layer_cls._b200c_orig_forward = layer_cls.forward
            layer_cls.forward = _layer_forward
        _wrap_model_init(getattr(mod, model_name))
