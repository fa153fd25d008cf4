"""Pipeline-stage hop over the P2P mailboxes.

vLLM 0.22 sends the tensors of an IntermediateTensors dict with torch.distributed.isend/irecv on
the NCCL group (vllm/distributed/parallel_state.py:871-926, 966-1050) and never asks the device
communicator.  This module routes CUDA tensors of same-node PP groups through
B200Communicator.send / recv (stream-ordered kernels writing the next stage's mailbox) and keeps
the reference protocol otherwise: pickled metadata over the CPU group
(parallel_state.py:674-735), TP slicing on the sender and a dim-0 all-gather on the receiver
(:913-916, :1014-1038).  The metadata of a stage boundary is static after the first step, so it
is cached and the per-step gloo handshake is skipped when both sides agree.
"""
from __future__ import annotations

from typing import Any

import torch

_APPLIED = False


def split_tensor_dict(tensor_dict: dict[str, Any]):
    """-> (metadata list, tensor list); tensors are replaced by (device type, dtype, size)."""
    from vllm.distributed.parallel_state import TensorMetadata
    meta, tensors = [], []
    for k, v in tensor_dict.items():
        if isinstance(v, torch.Tensor):
            meta.append((k, TensorMetadata(v.device.type, v.dtype, v.size())))
            tensors.append(v)
        else:
            meta.append((k, v))
    return meta, tensors


def _use_slices(group, key, numel, ag_group, ag_tensors) -> bool:
    return group._should_use_all_gather(key, numel, ag_group, ag_tensors)


def _b200(group):
    comm = group.device_communicator
    return comm if comm is not None and getattr(comm, "b200", None) is not None else None


def isend_tensor_dict(self, tensor_dict, dst=None, all_gather_group=None, all_gather_tensors=None):
    comm = _b200(self)
    if comm is None or self.world_size <= 1:
        return self._b200c_orig_isend(tensor_dict, dst, all_gather_group, all_gather_tensors)
    if dst is None:
        dst = (self.rank_in_group + 1) % self.world_size
    ag_size = 1 if all_gather_group is None else all_gather_group.world_size
    ag_rank = 0 if all_gather_group is None else all_gather_group.rank_in_group
    meta, tensors = split_tensor_dict(tensor_dict)
    self.send_object(meta, dst=dst)
    handles = []
    for (key, _), t in zip([m for m in meta if isinstance(tensor_dict[m[0]], torch.Tensor)], tensors):
        if t.numel() == 0:
            continue
        if _use_slices(self, key, t.numel(), all_gather_group, all_gather_tensors):
            t = t.reshape(ag_size, -1)[ag_rank]
        if t.is_cuda:
            tc = t.contiguous()
            if hasattr(comm, "isend") and comm.b200.can_p2p(tc):
                handles.append(comm.isend(tc, dst))  # side stream; never blocks the host
            else:
                comm.send(tc, dst)
        else:
            handles.append(torch.distributed.isend(t, dst=self.ranks[dst], group=self.cpu_group))
    return handles


def irecv_tensor_dict(self, src=None, all_gather_group=None, all_gather_tensors=None):
    comm = _b200(self)
    if comm is None or self.world_size <= 1:
        return self._b200c_orig_irecv(src, all_gather_group, all_gather_tensors)
    from vllm.distributed.parallel_state import TensorMetadata
    if src is None:
        src = (self.rank_in_group - 1) % self.world_size
    ag_size = 1 if all_gather_group is None else all_gather_group.world_size
    meta = self.recv_object(src=src)
    out: dict[str, Any] = {}
    handles, post = [], []
    for key, value in meta:
        if not isinstance(value, TensorMetadata):
            out[key] = value
            continue
        full_numel = 1
        for d in value.size:
            full_numel *= d
        if full_numel == 0:
            out[key] = torch.empty(value.size, dtype=value.dtype, device=value.device)
            continue
        sliced = _use_slices(self, key, full_numel, all_gather_group, all_gather_tensors)
        shape = torch.Size((full_numel // ag_size,)) if sliced else value.size
        if value.device == "cpu":
            t = torch.empty(shape, dtype=value.dtype, device="cpu")
            handles.append(torch.distributed.irecv(t, src=self.ranks[src], group=self.cpu_group))
        elif hasattr(comm, "irecv"):
            t, h = comm.irecv(shape, value.dtype, src)
            handles.append(h)
        else:
            t = comm.recv(shape, value.dtype, src)
        out[key] = t
        if sliced:
            def _join(key=key, t=t, shape=tuple(value.size), g=all_gather_group):
                out[key] = g.all_gather(t, dim=0).reshape(shape)
            post.append(_join)
    return out, handles, post


def apply() -> None:
    global _APPLIED
    if _APPLIED:
        return
    _APPLIED = True
    from vllm.distributed.parallel_state import GroupCoordinator
    if not hasattr(GroupCoordinator, "_b200c_orig_isend"):
        GroupCoordinator._b200c_orig_isend = GroupCoordinator.isend_tensor_dict
        GroupCoordinator._b200c_orig_irecv = GroupCoordinator.irecv_tensor_dict
        GroupCoordinator.isend_tensor_dict = isend_tensor_dict
        GroupCoordinator.irecv_tensor_dict = irecv_tensor_dict
