"""b200c as a provider of vLLM's IR ops rms_norm / fused_add_rms_norm (the TP=1 seam).

At TP=1 there is no communicator and no all-reduce (parallel_state.py:530-531): the only piece
of the hot path that runs is the residual-add + RMSNorm, which vLLM dispatches through
`vllm.ir.ops.fused_add_rms_norm` to the provider first in `kernel_config.ir_op_priority`
(vllm/ir/op.py:321-360).  Stock vLLM registers its `_C` kernels as provider "vllm_c"
(vllm/kernels/vllm_c.py:22-91) and CudaPlatform puts it first whenever Inductor is not generating
the code (vllm/platforms/cuda.py:574-593).  This module registers libb200c's norm kernels as
provider "b200c" with the same signatures, and puts it in front of "vllm_c" in exactly those cases.
"""
from __future__ import annotations

import os

import torch
from torch import Tensor

PROVIDER = "b200c"
_PATCHED_ATTR = "_b200c_ir_priority_patched"


_DTYPES = (torch.bfloat16, torch.float16, torch.float32)


def _takes(x: Tensor, weight: Tensor | None) -> bool:
    # on the dispatch hot path (vllm/ir/op.py:321): cheapest tests first
    if not x.is_cuda or x.dtype not in _DTYPES or not x.is_contiguous():
        return False
    row_bytes = x.shape[-1] * x.element_size() if x.dim() else 0
    return (row_bytes % 16 == 0 and 0 < row_bytes <= 2048 * 16 and x.numel() > 0
            and x.data_ptr() % 16 == 0
            and (weight is None or (weight.dtype == x.dtype and weight.is_contiguous()
                                    and weight.data_ptr() % 16 == 0)))


def _ones_like_row(x: Tensor) -> Tensor:
    return torch.ones(x.shape[-1], device=x.device, dtype=x.dtype)


def register(supported: bool | None = None) -> bool:
    """Register the two implementations (idempotent).  -> True if they are registered."""
    from vllm import ir
    if PROVIDER in ir.ops.rms_norm.impls:
        return True
    if supported is None:
        from vllm.platforms import current_platform
        supported = bool(current_platform.is_cuda())
    from . import comm as bc

    def rms_args(x, weight, epsilon, variance_size=None):
        return variance_size is None and _takes(x, weight)

    @ir.ops.rms_norm.register_impl(PROVIDER, supports_args=rms_args, supported=supported)
    def rms_norm(x: Tensor, weight: Tensor | None, epsilon: float,
                 variance_size: int | None = None) -> Tensor:
        if weight is None:
            weight = _ones_like_row(x)
        return bc.rms_norm(x, weight, epsilon)

    def add_args(x, x_residual, weight, epsilon, variance_size=None):
        return (variance_size is None and _takes(x, weight) and x_residual.is_cuda
                and x_residual.dtype == x.dtype and x_residual.shape == x.shape
                and x_residual.is_contiguous() and x_residual.data_ptr() % 16 == 0)

    @ir.ops.fused_add_rms_norm.register_impl(PROVIDER, supports_args=add_args, supported=supported,
                                             inplace=True)
    def fused_add_rms_norm(x: Tensor, x_residual: Tensor, weight: Tensor | None, epsilon: float,
                           variance_size: int | None = None) -> tuple[Tensor, Tensor]:
        if weight is None:
            weight = _ones_like_row(x)
        # in place, like the provider it replaces: x becomes the normalised output and
        # x_residual the updated residual stream
        return bc.fused_add_rms_norm_(x, x_residual, weight, epsilon)

    return True


def prioritise(platform_cls) -> None:
    """Wrap platform_cls.get_default_ir_op_priority so that "b200c" precedes "vllm_c" wherever
    vLLM would have used its own C kernels (no Inductor code generation).  With
    B200C_NORM_PROVIDER=always it is put first unconditionally, =off leaves vLLM's choice."""
    mode = os.environ.get("B200C_NORM_PROVIDER", "auto")
    if mode == "off" or getattr(platform_cls, _PATCHED_ATTR, False):
        return
    orig = platform_cls.get_default_ir_op_priority.__func__

    def patched(cls, vllm_config):
        prio = orig(cls, vllm_config)
        for name in ("rms_norm", "fused_add_rms_norm"):
            lst = list(getattr(prio, name))
            if PROVIDER in lst:
                continue
            if mode == "always":
                lst.insert(0, PROVIDER)
            elif "vllm_c" in lst:
                lst.insert(lst.index("vllm_c"), PROVIDER)
            else:
                continue
            setattr(prio, name, lst)
        return prio

    platform_cls.get_default_ir_op_priority = classmethod(patched)
    setattr(platform_cls, _PATCHED_ATTR, True)
