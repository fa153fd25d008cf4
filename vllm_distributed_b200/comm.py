"""Torch-facing host objects over the C ABI: one `B200Comm` per (process, group).

PyTorch is plumbing here (device memory, streams, the gloo group used once to swap IPC
handles); the arithmetic and the exchange run in libb200c.so.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import byref, c_longlong, c_size_t, c_void_p

import torch

from . import _cabi as C

DTYPE_CODE = {torch.bfloat16: C.BF16, torch.float16: C.F16, torch.float32: C.F32}

# Arena defaults, sized for B200's 180 GB: 128 MiB messages cover [8192 tok x 8192] bf16, i.e.
# vLLM's default max_num_batched_tokens of prefill at the 70B hidden size (with 64 MiB that
# point silently went to the stock path; three rotating regions of 2x that), 72 MiB covers the
# 70B lm_head gather at batch 256.
DEFAULT_MAX_BYTES = int(os.environ.get("B200C_MAX_BYTES", 128 << 20))
DEFAULT_AG_BYTES = int(os.environ.get("B200C_AG_BYTES", 72 << 20))
DEFAULT_P2P_BYTES = int(os.environ.get("B200C_P2P_BYTES", 16 << 20))


try:  # raw handle of the current stream of a device: about 10x cheaper than the Stream object
    _raw_stream = torch._C._cuda_getCurrentRawStream
except AttributeError:  # pragma: no cover -- older / CPU-only torch builds
    def _raw_stream(index: int) -> int:
        return torch.cuda.current_stream(index).cuda_stream


def _stream() -> int:
    """The current stream of the current device (what every entry point launches on)."""
    return _raw_stream(torch.cuda.current_device())

_NORM_FNS: list = []


def _norm_fns():
    """(fused_add_rms_norm, rms_norm) entry points, looked up once (eager decode is host-bound:
    65 norm calls per step at TP=1)."""
    if not _NORM_FNS:
        lib = C.lib()
        _NORM_FNS.extend([lib.b200c_fused_add_rms_norm, lib.b200c_rms_norm])
    return _NORM_FNS


def _ptr(t: torch.Tensor | None) -> int:
    return 0 if t is None else t.data_ptr()


def arena_bytes(world: int, max_bytes: int, ag_bytes: int, p2p_bytes: int) -> int:
    out = c_size_t(0)
    C.check(C.lib().b200c_arena_bytes(world, max_bytes, ag_bytes, p2p_bytes, byref(out)))
    return out.value


def alloc_shared(nbytes: int) -> tuple[int, bytes]:
    """cudaMalloc on the current device + its IPC handle."""
    ptr = c_void_p(0)
    handle = C.Handle()
    C.check(C.lib().b200c_alloc_shared(nbytes, byref(ptr), handle))
    return ptr.value, bytes(handle)


def open_handle(handle: bytes) -> int:
    ptr = c_void_p(0)
    h = C.Handle.from_buffer_copy(handle)
    C.check(C.lib().b200c_open_handle(h, byref(ptr)))
    return ptr.value


def assert_peer_access(cpu_group, device: torch.device) -> None:
    """Every pair of the group's GPUs must be able to map each other's memory, or the first
    push would fault.  The reference checks the same before it builds its IPC buffers
    (vllm/distributed/device_communicators/custom_all_reduce.py:151-163) and then falls back to
    NCCL; here the caller (B200Communicator) catches PeerAccessError and does the same."""
    import torch.distributed as dist
    if device.type != "cuda":
        return
    world = dist.get_world_size(cpu_group)
    mine = device.index if device.index is not None else torch.cuda.current_device()
    idx: list = [None] * world
    dist.all_gather_object(idx, mine, group=cpu_group)
    bad = [j for j in idx if j != mine and not torch.cuda.can_device_access_peer(mine, j)]
    flags: list = [None] * world
    dist.all_gather_object(flags, bad, group=cpu_group)
    if any(flags):
        raise PeerAccessError(f"no peer access between GPUs {[(idx[r], b) for r, b in enumerate(flags) if b]}")


class PeerAccessError(RuntimeError):
    pass


# ---- standalone norm ops (TP=1 and the unfused path; no communicator state) ----------------

def _rows(x: torch.Tensor) -> tuple[int, int]:
    hidden = x.shape[-1]
    return x.numel() // hidden if hidden else 0, hidden


def fused_add_rms_norm_(x: torch.Tensor, residual: torch.Tensor, gamma: torch.Tensor,
                        eps: float) -> tuple[torch.Tensor, torch.Tensor]:
    """In place: residual += x; x = rmsnorm(residual) * gamma.  (vLLM _C.fused_add_rms_norm)"""
    assert x.is_cuda and x.is_contiguous() and residual.is_contiguous()
    hidden = x.shape[-1]
    rc = _norm_fns()[0](x.data_ptr(), residual.data_ptr(), gamma.data_ptr(), eps,
                        x.numel() // hidden if hidden else 0, hidden, DTYPE_CODE[x.dtype],
                        _raw_stream(x.device.index))
    if rc:
        C.check(rc)
    return x, residual


def rms_norm(x: torch.Tensor, gamma: torch.Tensor, eps: float,
             out: torch.Tensor | None = None) -> torch.Tensor:
    assert x.is_cuda and x.is_contiguous()
    out = torch.empty_like(x) if out is None else out
    hidden = x.shape[-1]
    rc = _norm_fns()[1](out.data_ptr(), x.data_ptr(), gamma.data_ptr(), eps,
                        x.numel() // hidden if hidden else 0, hidden, DTYPE_CODE[x.dtype],
                        _raw_stream(x.device.index))
    if rc:
        C.check(rc)
    return out


def choose_thresholds(table: dict[int, dict[str, float]]) -> dict[str, int]:
    """From measured us per call {bytes: {"oneshot": t, "twoshot": t[, "nvls": t]}} pick
    `oneshot_max_bytes` = the largest size up to which one-shot never lost to two-shot (scanning
    upwards, so one noisy point higher up cannot extend it), and `nvls_min_bytes` = the smallest
    size from which the in-switch variant wins at every larger measured size (-1 if it never does).
    Pure function of its input: every rank feeds it the same max-over-ranks table and gets the
    same answer, which the protocol requires (all ranks must run the same variant)."""
    sizes = sorted(table)
    oneshot_max = 0
    for sz in sizes:
        t = table[sz]
        if "oneshot" in t and t["oneshot"] <= t["twoshot"]:
            oneshot_max = sz
        else:
            break
    nvls_min = -1
    for sz in reversed(sizes):
        t = table[sz]
        best_push = min(v for k, v in t.items() if k != "nvls")
        if "nvls" in t and t["nvls"] < best_push:
            nvls_min = sz
        else:
            break
    return {"oneshot_max_bytes": oneshot_max, "nvls_min_bytes": nvls_min}


class B200Comm:
    """One rank of a TP (or PP) group.  Owns its arena and the peers' IPC mappings."""

    def __init__(self, rank: int, world: int, arena_ptrs: list[int] | None, max_bytes: int,
                 ag_bytes: int, p2p_bytes: int, device: torch.device, owned_arena: int = 0,
                 opened: list[int] | None = None):
        self.rank, self.world, self.device = rank, world, device
        self.max_bytes, self.ag_bytes, self.p2p_bytes = max_bytes, ag_bytes, p2p_bytes
        self._owned_arena = owned_arena
        self._opened = opened or []
        self.nvls = False          # True once a multicast buffer is attached
        self._nvls_buf = None      # keeps the symmetric allocation alive
        ctx = c_void_p(0)
        arr = C.ptr_array(arena_ptrs) if arena_ptrs else None
        with torch.cuda.device(device):
            C.check(C.lib().b200c_init(rank, world, arr, max_bytes, ag_bytes, p2p_bytes, byref(ctx)))
        self.ctx = ctx

    # ---- construction over a torch.distributed CPU group (one process per GPU) ----
    @classmethod
    def from_process_group(cls, cpu_group, device: torch.device,
                           max_bytes: int = DEFAULT_MAX_BYTES, ag_bytes: int = DEFAULT_AG_BYTES,
                           p2p_bytes: int = DEFAULT_P2P_BYTES, device_group=None) -> "B200Comm":
        """cpu_group: a gloo group (handle exchange, barriers).  device_group: the NCCL group of
        the same ranks; when given (and B200C_NVLS != 0) a symmetric-memory buffer is
        rendezvoused over it and, if the fabric offers multicast, attached for the NVLS variant."""
        import torch.distributed as dist
        rank = dist.get_rank(cpu_group)
        world = dist.get_world_size(cpu_group)
        if world == 1:
            return cls(0, 1, None, 0, 0, 0, device)
        assert_peer_access(cpu_group, device)
        nbytes = arena_bytes(world, max_bytes, ag_bytes, p2p_bytes)
        with torch.cuda.device(device):
            mine, handle = alloc_shared(nbytes)
            handles: list = [None] * world
            dist.all_gather_object(handles, handle, group=cpu_group)
            ptrs, opened = [], []
            for r in range(world):
                if r == rank:
                    ptrs.append(mine)
                else:
                    p = open_handle(handles[r])
                    ptrs.append(p)
                    opened.append(p)
            comm = cls(rank, world, ptrs, max_bytes, ag_bytes, p2p_bytes, device,
                       owned_arena=mine, opened=opened)
        if device_group is not None and os.environ.get("B200C_NVLS", "1") != "0":
            comm.try_attach_nvls(cpu_group, device_group)
        # every arena must be armed before any rank pushes into it
        dist.barrier(group=cpu_group)
        return comm

    def try_attach_nvls(self, cpu_group, device_group) -> bool:
        """Allocate + rendezvous a symmetric buffer (torch plumbing) and hand its unicast and
        multicast mappings to the library.  All-or-nothing across the group: if any rank has no
        multicast pointer (no NVSwitch, driver without fabric support) nobody attaches."""
        import torch.distributed as dist
        need = c_size_t(0)
        C.check(C.lib().b200c_nvls_bytes(self.max_bytes, byref(need)))
        mc = uc = 0
        buf = None
        why = ""
        # step 1, local: allocate.  The ranks agree on the outcome BEFORE the rendezvous, which is
        # a collective: a rank that failed here must not leave its peers waiting in it.
        try:
            import warnings
            import torch.distributed._symmetric_memory as symm_mem
            name = device_group.group_name
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                try:
                    symm_mem.enable_symm_mem_for_group(name)
                except Exception:  # noqa: BLE001 -- no-op or deprecated in newer torch
                    pass
            with torch.cuda.device(self.device):
                buf = symm_mem.empty(need.value, dtype=torch.uint8, device=self.device)
        except Exception as e:  # noqa: BLE001 -- the NVLS variant is optional
            why = f"{type(e).__name__}: {e}"
        allocated: list = [None] * self.world
        dist.all_gather_object(allocated, buf is not None, group=cpu_group)
        if not all(allocated):
            self.nvls_reason = why or "symmetric-memory allocation failed on some rank"
            return False
        # step 2, collective: rendezvous; a fabric without multicast answers with a null pointer
        try:
            with torch.cuda.device(self.device):
                hdl = symm_mem.rendezvous(buf, name)
            mc, uc = int(hdl.multicast_ptr), int(buf.data_ptr())
        except Exception as e:  # noqa: BLE001
            why = f"{type(e).__name__}: {e}"
        oks: list = [None] * self.world
        dist.all_gather_object(oks, bool(mc), group=cpu_group)
        if not all(oks):
            self.nvls_reason = why or "no multicast pointer on some rank"
            return False
        with torch.cuda.device(self.device):
            C.check(C.lib().b200c_attach_nvls(self.ctx, uc, mc, need.value))
        self._nvls_buf = buf
        self.nvls = True
        v = os.environ.get("B200C_NVLS_MIN_BYTES")
        if v is not None:
            self.set_tuning(C.TUNE_NVLS_MIN_BYTES, int(v))
        return True

    # ---- tuning ----
    def set_tuning(self, key: int, value: int) -> None:
        C.check(C.lib().b200c_set_tuning(self.ctx, key, int(value)))

    def get_tuning(self, key: int) -> int:
        v = c_longlong(0)
        C.check(C.lib().b200c_get_tuning(self.ctx, key, byref(v)))
        return v.value

    def check_health(self) -> None:
        C.check(C.lib().b200c_query_error(self.ctx))

    def calibrate(self, cpu_group, hidden: int = 8192, iters: int = 12,
                  sizes: tuple[int, ...] = (64 << 10, 256 << 10, 512 << 10, 1 << 20, 2 << 20, 4 << 20)) -> dict:
        """Measure the fused call per variant at a few sizes ON THIS BOX and set the AUTO
        thresholds from the result (SURVEY 8d "variant selection": a measured table written at
        init, not constants tuned on someone else's machine -- the compiled-in defaults come from
        one 8 x B200 box).  Collective: every rank of the group calls it at the same point.
        About 0.2 s.  NVLS sizes (32 / 64 MiB) are added when a multicast buffer is attached
        and the arena is large enough."""
        import torch.distributed as dist
        dev = self.device
        sizes = tuple(s for s in sizes if s <= self.max_bytes)
        if self.nvls:
            sizes += tuple(s for s in (32 << 20, 64 << 20) if s <= self.max_bytes)
        gamma = torch.ones(hidden, dtype=torch.bfloat16, device=dev)
        table: dict[int, dict[str, float]] = {}
        for sz in sizes:
            tokens = max(1, sz // (hidden * 2))
            x = torch.zeros(tokens, hidden, dtype=torch.bfloat16, device=dev)
            res = torch.zeros_like(x)
            out = torch.empty_like(x)
            names = [("twoshot", C.TWOSHOT)]
            if x.numel() * 2 * self.world <= 2 * self.max_bytes and sz <= (4 << 20):
                names.append(("oneshot", C.ONESHOT))
            if self.nvls and sz >= (4 << 20):
                names.append(("nvls", C.NVLS))
            row = {}
            for name, code in names:
                for _ in range(3):
                    self.all_reduce_add_rmsnorm_into(x, res, gamma, 1e-5, out, code)
                torch.cuda.synchronize(dev)
                dist.barrier(group=cpu_group)
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(iters):
                    self.all_reduce_add_rmsnorm_into(x, res, gamma, 1e-5, out, code)
                b.record()
                torch.cuda.synchronize(dev)
                t = torch.tensor([a.elapsed_time(b) * 1e3 / iters], dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX, group=cpu_group)   # same number on every rank
                row[name] = float(t.item())
            table[x.numel() * 2] = row
        chosen = choose_thresholds(table)
        if chosen["oneshot_max_bytes"] > 0:
            self.set_tuning(C.TUNE_ONESHOT_MAX_BYTES, chosen["oneshot_max_bytes"])
        if self.nvls:
            self.set_tuning(C.TUNE_NVLS_MIN_BYTES, chosen["nvls_min_bytes"])
        self.variant_table = {"us_per_call": table, "chosen": chosen}
        return self.variant_table

    # ---- predicates the communicator uses to decide between this path and NCCL ----
    def can_all_reduce(self, x: torch.Tensor) -> bool:
        # (every size up to max_bytes fits both variants for every group size: the arena
        # regions carry the chunk-rounding slack, see make_geom in csrc/b200c.cu)
        nbytes = x.numel() * x.element_size()
        return (x.is_cuda and x.dtype in DTYPE_CODE and x.is_contiguous() and nbytes % 16 == 0
                and 0 < nbytes <= self.max_bytes and x.data_ptr() % 16 == 0)

    def can_fuse_norm(self, x: torch.Tensor) -> bool:
        hidden = x.shape[-1]
        return (self.can_all_reduce(x) and (hidden * x.element_size()) % 16 == 0
                and hidden * x.element_size() // 16 <= 2048)

    def can_all_gather_lastdim(self, x: torch.Tensor) -> bool:
        row_bytes = x.shape[-1] * x.element_size()
        vec = 16 if row_bytes % 16 == 0 else 8 if row_bytes % 8 == 0 else 4
        return (x.is_cuda and x.is_contiguous() and row_bytes % 4 == 0 and x.numel() > 0
                and x.numel() * x.element_size() * self.world <= self.ag_bytes
                and x.data_ptr() % vec == 0)

    def can_p2p(self, t: torch.Tensor) -> bool:
        nbytes = t.numel() * t.element_size()
        return t.is_cuda and t.is_contiguous() and nbytes % 16 == 0 and t.data_ptr() % 16 == 0

    # ---- the hot path ----
    def all_reduce(self, x: torch.Tensor, variant: int = C.AUTO,
                   out: torch.Tensor | None = None) -> torch.Tensor:
        out = torch.empty_like(x) if out is None else out
        C.check(C.lib().b200c_all_reduce(self.ctx, x.data_ptr(), out.data_ptr(), x.numel(),
                                         DTYPE_CODE[x.dtype], variant, _stream()))
        return out

    def all_reduce_add_rmsnorm(self, x: torch.Tensor, residual: torch.Tensor | None,
                               gamma: torch.Tensor, eps: float, variant: int = C.AUTO,
                               inplace: bool = False) -> tuple[torch.Tensor, torch.Tensor | None]:
        """-> (norm_out, residual_out).  inplace: norm_out aliases x, residual_out aliases residual."""
        tokens, hidden = _rows(x)
        norm_out = x if inplace else torch.empty_like(x)
        res_out = None if residual is None else (residual if inplace else torch.empty_like(residual))
        C.check(C.lib().b200c_all_reduce_add_rmsnorm(
            self.ctx, x.data_ptr(), _ptr(residual), _ptr(res_out), norm_out.data_ptr(),
            gamma.data_ptr(), float(eps), tokens, hidden, DTYPE_CODE[x.dtype], variant, _stream()))
        return norm_out, res_out

    def all_reduce_add_rmsnorm_into(self, x: torch.Tensor, residual: torch.Tensor,
                                    gamma: torch.Tensor, eps: float, norm_out: torch.Tensor,
                                    variant: int = C.AUTO) -> None:
        """residual updated in place (the residual stream), normalised output into norm_out."""
        tokens, hidden = _rows(x)
        C.check(C.lib().b200c_all_reduce_add_rmsnorm(
            self.ctx, x.data_ptr(), residual.data_ptr(), residual.data_ptr(), norm_out.data_ptr(),
            gamma.data_ptr(), float(eps), tokens, hidden, DTYPE_CODE[x.dtype], variant, _stream()))

    def all_gather_lastdim(self, x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
        rows, cols = _rows(x)
        if out is None:
            out = torch.empty(x.shape[:-1] + (cols * self.world,), dtype=x.dtype, device=x.device)
        C.check(C.lib().b200c_all_gather_lastdim(self.ctx, x.data_ptr(), out.data_ptr(), rows,
                                                 cols * x.element_size(), _stream()))
        return out

    def send(self, t: torch.Tensor, dst: int) -> None:
        nbytes = t.numel() * t.element_size()
        base = t.data_ptr()
        for off in range(0, nbytes, self.p2p_bytes):
            n = min(self.p2p_bytes, nbytes - off)
            C.check(C.lib().b200c_send(self.ctx, dst, base + off, n, _stream()))

    def recv(self, t: torch.Tensor, src: int) -> torch.Tensor:
        nbytes = t.numel() * t.element_size()
        base = t.data_ptr()
        for off in range(0, nbytes, self.p2p_bytes):
            n = min(self.p2p_bytes, nbytes - off)
            C.check(C.lib().b200c_recv(self.ctx, src, base + off, n, _stream()))
        return t

    def close(self) -> None:
        if self.ctx:
            with torch.cuda.device(self.device):
                C.lib().b200c_destroy(self.ctx)
                self.ctx = c_void_p(0)
                self._nvls_buf = None
                self.nvls = False
                for p in self._opened:
                    C.lib().b200c_close_handle(p)
                self._opened = []
                if self._owned_arena:
                    C.lib().b200c_free_shared(self._owned_arena)
                    self._owned_arena = 0

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
