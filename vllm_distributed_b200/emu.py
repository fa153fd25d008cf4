"""Single-GPU emulation of an N-rank group (bring-up harness and the 1-GPU parity tests).

All ranks' arenas live on one device and ONE cooperative launch runs every rank's CTAs of
the same device code the real entry points run (see include/b200c.h, "single-GPU rank
emulation").  Kernels that wait on each other must never be separate launches on one GPU.
"""
from __future__ import annotations

from ctypes import c_void_p

import torch

from . import _cabi as C
from .comm import DTYPE_CODE, B200Comm, _ptr, _rows, _stream, alloc_shared, arena_bytes


class EmuGroup:
    def __init__(self, world: int, device: torch.device | str = "cuda:0",
                 max_bytes: int = 8 << 20, ag_bytes: int = 8 << 20, p2p_bytes: int = 4 << 20):
        self.world = world
        self.device = torch.device(device)
        nbytes = arena_bytes(world, max_bytes, ag_bytes, p2p_bytes)
        self._arenas = []
        with torch.cuda.device(self.device):
            for _ in range(world):
                p, _h = alloc_shared(nbytes)
                self._arenas.append(p)
        self.ranks = [B200Comm(r, world, self._arenas, max_bytes, ag_bytes, p2p_bytes, self.device)
                      for r in range(world)]
        self._ctxs = C.ptr_array([c.ctx.value for c in self.ranks])

    def set_tuning(self, key: int, value: int) -> None:
        for c in self.ranks:
            c.set_tuning(key, value)

    def check_health(self) -> None:
        for c in self.ranks:
            c.check_health()

    def all_reduce(self, xs: list[torch.Tensor], variant: int = C.AUTO,
                   outs: list[torch.Tensor] | None = None) -> list[torch.Tensor]:
        outs = [torch.empty_like(x) for x in xs] if outs is None else outs
        C.check(C.lib().b200c_emu_all_reduce(
            self._ctxs, self.world, C.ptr_array([x.data_ptr() for x in xs]),
            C.ptr_array([o.data_ptr() for o in outs]), xs[0].numel(), DTYPE_CODE[xs[0].dtype],
            variant, _stream()))
        return outs

    def all_reduce_add_rmsnorm(self, xs, residuals, gamma, eps, variant: int = C.AUTO,
                               inplace: bool = False):
        """-> (norm_outs, residual_outs), one per rank."""
        tokens, hidden = _rows(xs[0])
        norm_outs = list(xs) if inplace else [torch.empty_like(x) for x in xs]
        if residuals is None:
            residuals = [None] * self.world
            res_outs = [None] * self.world
        else:
            res_outs = list(residuals) if inplace else [torch.empty_like(r) for r in residuals]
        C.check(C.lib().b200c_emu_all_reduce_add_rmsnorm(
            self._ctxs, self.world, C.ptr_array([x.data_ptr() for x in xs]),
            C.ptr_array([_ptr(r) for r in residuals]), C.ptr_array([_ptr(r) for r in res_outs]),
            C.ptr_array([o.data_ptr() for o in norm_outs]), gamma.data_ptr(), float(eps), tokens,
            hidden, DTYPE_CODE[xs[0].dtype], variant, _stream()))
        return norm_outs, res_outs

    def all_gather_lastdim(self, xs: list[torch.Tensor]) -> list[torch.Tensor]:
        rows, cols = _rows(xs[0])
        outs = [torch.empty(x.shape[:-1] + (cols * self.world,), dtype=x.dtype, device=x.device)
                for x in xs]
        C.check(C.lib().b200c_emu_all_gather_lastdim(
            self._ctxs, self.world, C.ptr_array([x.data_ptr() for x in xs]),
            C.ptr_array([o.data_ptr() for o in outs]), rows, cols * xs[0].element_size(),
            _stream()))
        return outs

    def sendrecv(self, src: int, dst: int, t: torch.Tensor,
                 out: torch.Tensor | None = None) -> torch.Tensor:
        out = torch.empty_like(t) if out is None else out
        nbytes = t.numel() * t.element_size()
        step = self.ranks[0].p2p_bytes
        for off in range(0, nbytes, step):
            n = min(step, nbytes - off)
            C.check(C.lib().b200c_emu_sendrecv(self._ctxs, self.world, src, dst,
                                               t.data_ptr() + off, out.data_ptr() + off, n,
                                               _stream()))
        return out

    def close(self) -> None:
        for c in self.ranks:
            c.close()
        with torch.cuda.device(self.device):
            for p in self._arenas:
                C.lib().b200c_free_shared(p)
        self._arenas = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
