"""B200Communicator: drop-in for vLLM's CudaCommunicator on the TP / PP hot path.

Mirrors the interface GroupCoordinator drives (vllm/distributed/parallel_state.py:538-616,
1061-1091): all_reduce, all_gather, send, recv, destroy -- same names, argument meaning and
error behaviour as vllm/distributed/device_communicators/cuda_communicator.py:241-420 and
base_device_communicator.py:184-300.  It subclasses CudaCommunicator because
GroupCoordinator.graph_capture asserts the type (parallel_state.py:489-496); everything it
does not take (non-contiguous or odd-sized tensors, messages beyond the arena, groups that
span hosts, groups that are neither TP nor PP) goes to the parent class unchanged.
"""
from __future__ import annotations

import os

import torch
from vllm.distributed.device_communicators.cuda_communicator import CudaCommunicator
from vllm.logger import init_logger

from . import _cabi as C
from .comm import (DEFAULT_AG_BYTES, DEFAULT_MAX_BYTES, DEFAULT_P2P_BYTES, B200Comm,
                   PeerAccessError)

logger = init_logger("vllm.b200c.communicator")  # under vLLM's logging tree so INFO is shown


def _group_kind(unique_name: str) -> str:
    return unique_name.split(":")[0]


class StreamHandle:
    """vLLM's Handle protocol (parallel_state.py:73-78) over a CUDA stream: wait() makes the
    current stream wait for everything enqueued so far on the side stream -- no host block."""

    def __init__(self, stream: "torch.cuda.Stream", wait_needed: bool):
        self.event = torch.cuda.Event()
        self.event.record(stream)
        self.wait_needed = wait_needed

    def is_completed(self) -> bool:
        return self.event.query()

    def wait(self) -> None:
        if self.wait_needed:
            torch.cuda.current_stream().wait_event(self.event)


class B200Communicator(CudaCommunicator):
    def __init__(self, cpu_group, device=None, device_group=None, unique_name: str = "", **kw):
        super().__init__(cpu_group, device, device_group, unique_name, **kw)
        self.b200: B200Comm | None = None
        C.lib()  # a missing libb200c.so is an error, never a silent NCCL run
        kind = _group_kind(unique_name)
        if self.world_size < 2 or self.world_size > C.MAX_RANKS or kind not in ("tp", "pp"):
            return
        if os.environ.get("B200C_DISABLE", "0") == "1":
            return
        from vllm.distributed.parallel_state import in_the_same_node_as
        if not all(in_the_same_node_as(cpu_group, source_rank=0)):
            # the reference launcher keeps a TP group on one host (launch.py:213-240); PP hops
            # between hosts stay on NCCL exactly as in the reference
            logger.info("b200c: group %s spans hosts, using the stock communicator", unique_name)
            return
        if kind == "tp":
            sizes = dict(max_bytes=DEFAULT_MAX_BYTES, ag_bytes=DEFAULT_AG_BYTES, p2p_bytes=1 << 20)
        else:
            sizes = dict(max_bytes=1 << 20, ag_bytes=1 << 20, p2p_bytes=DEFAULT_P2P_BYTES)
        try:
            self.b200 = B200Comm.from_process_group(
                cpu_group, self.device, device_group=device_group if kind == "tp" else None, **sizes)
        except PeerAccessError as e:
            # same decision as the reference's custom all-reduce (custom_all_reduce.py:151-163):
            # without full peer access the stock communicator serves the group
            logger.warning("b200c: %s; group %s stays on the stock communicator", e, unique_name)
            return
        # A wait on a peer that never answers: the reference's barrier would hang forever
        # (vllm_custom_all_reduce.cuh:190-211).  Here the wait is bounded by minutes, not seconds
        # (one rank may legitimately be late: a JIT compile, a host stall, a long upstream PP
        # stage), and when it does expire the kernel traps, so the worker dies loudly -- the
        # launcher turns that into its fatal-disconnect path (/root/reference/src/launch.py:130-144)
        # -- instead of serving tokens computed from garbage.
        self.b200.set_tuning(C.TUNE_SPIN_TIMEOUT_MS, int(os.environ.get("B200C_SPIN_TIMEOUT_MS", 300000)))
        self.b200.set_tuning(C.TUNE_TRAP_ON_TIMEOUT, int(os.environ.get("B200C_TRAP_ON_TIMEOUT", 1)))
        if kind == "tp" and os.environ.get("B200C_CALIBRATE", "0") == "1":
            chosen = self.b200.calibrate(cpu_group)["chosen"]
            logger.info("b200c: AUTO thresholds measured on this box: %s", chosen)
        logger.info("b200c: %s group of %d ranks on NVLink arenas (%s)%s", kind, self.world_size,
                    ", ".join(f"{k}={v >> 20} MiB" for k, v in sizes.items()),
                    ", NVLS multicast attached" if self.b200.nvls else "")

    # ---- a1..a5: all-reduce -------------------------------------------------------------------
    def all_reduce(self, input_: torch.Tensor) -> torch.Tensor:
        b = self.b200
        if b is not None and b.can_all_reduce(input_):
            return b.all_reduce(input_)
        return super().all_reduce(input_)

    # ---- a6..a9: fused all-reduce + residual + RMSNorm ---------------------------------------
    def can_fuse_norm(self, x: torch.Tensor) -> bool:
        return self.b200 is not None and x.dim() >= 2 and self.b200.can_fuse_norm(x)

    def all_reduce_add_rmsnorm(self, x: torch.Tensor, residual: torch.Tensor | None,
                               weight: torch.Tensor, eps: float,
                               inplace: bool = True) -> tuple[torch.Tensor, torch.Tensor | None]:
        """(norm_out, residual_out) = fused_add_rms_norm(all_reduce(x), residual, weight, eps).
        With inplace=True x becomes norm_out and residual is updated, like the reference's
        flashinfer_trtllm_fused_allreduce_norm (allreduce_rms_fusion.py:121-212)."""
        if self.can_fuse_norm(x) and weight.dtype == x.dtype and weight.is_contiguous():
            return self.b200.all_reduce_add_rmsnorm(x, residual, weight, eps, inplace=inplace)
        # unfused path: the parent's all-reduce and this library's local norm kernels
        from .comm import fused_add_rms_norm_, rms_norm
        ar = self.all_reduce(x)
        if residual is None:
            return rms_norm(ar, weight, eps), None
        if not inplace:
            residual = residual.clone()
        fused_add_rms_norm_(ar, residual, weight, eps)
        return ar, residual

    # ---- a10..a11: lm_head logits gather ------------------------------------------------------
    def all_gather(self, input_: torch.Tensor, dim: int = -1) -> torch.Tensor:
        b = self.b200
        if dim < 0:
            dim += input_.dim()
        if (b is not None and input_.dim() >= 1 and dim == input_.dim() - 1
                and b.can_all_gather_lastdim(input_)):
            return b.all_gather_lastdim(input_)
        if b is not None and dim == 0 and input_.is_contiguous() and input_.numel() > 0:
            # dim-0 gather (the PP re-assembly, parallel_state.py:1014-1038) is the last-dim
            # gather of the flattened tensor viewed as one row
            flat = input_.reshape(1, -1)
            if b.can_all_gather_lastdim(flat):
                out = b.all_gather_lastdim(flat)
                return out.reshape((self.world_size * input_.shape[0],) + tuple(input_.shape[1:]))
        return super().all_gather(input_, dim)

    # ---- a12..a13: pipeline hop ------------------------------------------------------------------
    def send(self, tensor: torch.Tensor, dst: int | None = None) -> None:
        if dst is None:
            dst = (self.rank_in_group + 1) % self.world_size
        b = self.b200
        if (b is not None and tensor.device.type == self.device.type
                and (tensor.numel() * tensor.element_size()) % 16 == 0):
            # the receiver decides by the byte count alone (it allocates an aligned tensor), so
            # the sender must too: a strided or misaligned source is copied, not re-routed
            if not tensor.is_contiguous() or tensor.data_ptr() % 16:
                tensor = tensor.contiguous() if not tensor.is_contiguous() else tensor.clone()
            b.send(tensor, dst)
            return
        super().send(tensor, dst)

    def recv(self, size: torch.Size, dtype: torch.dtype, src: int | None = None) -> torch.Tensor:
        if src is None:
            src = (self.rank_in_group - 1) % self.world_size
        b = self.b200
        if b is not None:
            tensor = torch.empty(size, dtype=dtype, device=self.device)
            if b.can_p2p(tensor):
                return b.recv(tensor, src)
        return super().recv(size, dtype, src)

    # ---- asynchronous flavour used by the tensor-dict channel (pp_channel.py) -----------------
    # NCCL's isend/irecv run on NCCL's own stream, so the compute stream keeps preparing the
    # next step while a hop is in flight.  Same here: sends and receives are enqueued on two
    # side streams, ordered against the compute stream with events only.
    def _side_streams(self):
        if getattr(self, "_send_stream", None) is None:
            self._send_stream = torch.cuda.Stream(device=self.device)
            self._recv_stream = torch.cuda.Stream(device=self.device)
        return self._send_stream, self._recv_stream

    def isend(self, tensor: torch.Tensor, dst: int) -> "StreamHandle":
        send_stream, _ = self._side_streams()
        send_stream.wait_stream(torch.cuda.current_stream(self.device))  # tensor is produced first
        with torch.cuda.stream(send_stream):
            self.b200.send(tensor, dst)
        tensor.record_stream(send_stream)
        return StreamHandle(send_stream, wait_needed=False)

    def irecv(self, size: torch.Size, dtype: torch.dtype, src: int):
        _, recv_stream = self._side_streams()
        tensor = torch.empty(size, dtype=dtype, device=self.device)
        recv_stream.wait_stream(torch.cuda.current_stream(self.device))  # allocation is ordered first
        with torch.cuda.stream(recv_stream):
            self.b200.recv(tensor, src)
        tensor.record_stream(recv_stream)
        return tensor, StreamHandle(recv_stream, wait_needed=True)

    def check_health(self) -> None:
        """Raises B200CError if a device-side wait on a peer ever timed out (host-synchronous).
        The reference's barrier would have hung instead (vllm_custom_all_reduce.cuh:190-211)."""
        if self.b200 is not None:
            self.b200.check_health()

    def destroy(self):
        if self.b200 is not None:
            try:
                self.b200.check_health()
            except Exception as e:  # noqa: BLE001 -- report, then still release the arenas
                logger.error("b200c: %s", e)
            self.b200.close()
            self.b200 = None
        super().destroy()
