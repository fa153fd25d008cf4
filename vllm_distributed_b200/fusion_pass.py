"""Model-agnostic fusion under torch.compile: a post-grad Inductor pass that rewrites

    ar        = torch.ops.vllm.all_reduce(x, group)
    rms, res  = vllm_ir.fused_add_rms_norm(ar, residual, weight, eps)

into ONE call of torch.ops.vllm.b200c_allreduce_add_rmsnorm (in place on x and residual), for
every dense model whose decoder layers produce that pair -- the role of
AllReduceFusedAddRMSNormPattern in the reference (vllm/compilation/passes/fusion/
allreduce_rms_fusion.py:347-416; `maybe_inplace` calls have already been functionalised to the
default overload by then, compilation/passes/ir/inplace_functionalization.py).  The layer patch in
model_patch.py stays the eager path (and covers Llama/Qwen2 without compilation); where it is
active the graph already holds the fused op and this pass finds nothing to do.

Hooks: (i) `CompilationConfig(inductor_passes={"post_grad_custom_post_pass":
"vllm_distributed_b200.b200c_fusion_pass"})` -- the documented seam (config/compilation.py:582-587,
925-940); (ii) with B200C_FUSION_PASS=1 plugin.register() appends the pass to PostGradPassManager
itself, so that a user of the plugin does not have to edit the config (opt-in: the rewrite is
pinned on hand-built graphs but has not been run end to end on a non-Llama model yet).  Set
`pass_config.fuse_allreduce_rms=False` so that FlashInfer's pass does not claim the same pairs first.
"""
from __future__ import annotations

import operator

import torch
from torch import fx


def _is(node: fx.Node, packet_name: str, ns: str) -> bool:
    t = node.target
    return (node.op == "call_function" and isinstance(t, torch._ops.OpOverload)
            and t.namespace == ns and t._schema.name.split("::")[-1] == packet_name)


def _arg(node: fx.Node, pos: int, name: str, default=None):
    if len(node.args) > pos:
        return node.args[pos]
    return node.kwargs.get(name, default)


def fuse_allreduce_rmsnorm(graph: fx.Graph) -> int:
    """Rewrite every fusable pair in `graph`; returns how many were rewritten."""
    from torch._higher_order_ops.auto_functionalize import auto_functionalized
    fused_op = torch.ops.vllm.b200c_allreduce_add_rmsnorm.default
    count = 0
    for ar in list(graph.nodes):
        if not _is(ar, "all_reduce", "vllm") or len(ar.users) != 1:
            continue
        norm = next(iter(ar.users))
        if not _is(norm, "fused_add_rms_norm", "vllm_ir") or norm.target._overloadname != "default":
            continue
        x, group_name = _arg(ar, 0, "tensor"), _arg(ar, 1, "group_name")
        if _arg(norm, 0, "x") is not ar or not isinstance(group_name, str):
            continue
        residual, weight = _arg(norm, 1, "x_residual"), _arg(norm, 2, "weight")
        eps, var_size = _arg(norm, 3, "epsilon"), _arg(norm, 4, "variance_size")
        if weight is None or var_size is not None or residual is ar or not isinstance(eps, float):
            continue
        users = list(norm.users)
        if not all(u.op == "call_function" and u.target is operator.getitem for u in users):
            continue
        with graph.inserting_before(norm):
            fused = graph.call_function(auto_functionalized, (fused_op,), dict(
                allreduce_in=x, residual=residual, weight=weight, eps=eps, group_name=group_name))
            # auto_functionalized returns (op result, *new values of the mutated arguments)
            new_out = {0: graph.call_function(operator.getitem, (fused, 1)),   # normalised output
                       1: graph.call_function(operator.getitem, (fused, 2))}   # residual stream
        for u in users:
            u.replace_all_uses_with(new_out[u.args[1]])
            graph.erase_node(u)
        graph.erase_node(norm)
        graph.erase_node(ar)
        count += 1
    if count:
        graph.eliminate_dead_code()
    return count


def _make_pass():
    from vllm.compilation.passes.inductor_pass import InductorPass

    class B200CAllReduceNormFusionPass(InductorPass):
        """InductorPass wrapper (uuid = hash of this file, so the compile cache notices changes)."""

        def __init__(self):
            self.matched = 0

        def __call__(self, graph: fx.Graph) -> None:
            from vllm.distributed.parallel_state import get_tp_group
            tp = get_tp_group()
            comm = tp.device_communicator
            if tp.world_size < 2 or comm is None or getattr(comm, "b200", None) is None:
                return
            self.matched += fuse_allreduce_rmsnorm(graph)

        def uuid(self):
            return InductorPass.hash_source(fuse_allreduce_rmsnorm, _make_pass)

    return B200CAllReduceNormFusionPass()


class _LazyPass:
    """What `inductor_passes` resolves to (`__import__(pkg).__dict__[name]` wants an attribute of
    the top-level package): a callable that builds the real pass on first use."""

    def __init__(self):
        self._impl = None

    def __call__(self, graph: fx.Graph) -> None:
        if self._impl is None:
            self._impl = _make_pass()
        self._impl(graph)


def install_in_pass_manager() -> None:
    """plugin.register(): append the pass whenever vLLM configures its post-grad pass manager."""
    from vllm.compilation.passes.pass_manager import PostGradPassManager
    if getattr(PostGradPassManager.configure, "_b200c_wrapped", False):
        return
    orig = PostGradPassManager.configure

    def configure(self, config):
        orig(self, config)
        try:
            if not any(type(p).__name__ == "B200CAllReduceNormFusionPass" for p in self.passes):
                self.add(_make_pass())
        except Exception:  # noqa: BLE001 -- an optimisation must not stop the engine from compiling
            import logging
            logging.getLogger(__name__).exception("b200c: fusion pass not installed")
    configure._b200c_wrapped = True
    PostGradPassManager.configure = configure
