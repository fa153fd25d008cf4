"""Pipeline-stage hop over the P2P mailboxes.

vLLM 0.22 sends the tensors of an IntermediateTensors dict with torch.distributed.isend/irecv on
the NCCL group (vllm/distributed/parallel_state.py:871-926, 966-1050) and never asks the device
communicator.  This module routes CUDA tensors of same-node PP groups through
B200Communicator.send / recv (stream-ordered kernels writing the next stage's mailbox) and keeps
the reference protocol otherwise: TP slicing on the sender and a dim-0 all-gather on the receiver
(:913-916, :1014-1038).

Metadata fast path.  The reference pickles the metadata list and pushes it through
two blocking gloo messages per hop per step (send_object, parallel_state.py:674-735).  Here only
the FIRST message of a (sender, receiver) pair travels that way; it also carries the name of a
shared-memory ring the sender has created.  From then on the metadata goes through that ring
(same-node processes: a store and a spin on a cache line instead of a TCP round trip), and when
it is byte-identical to the previous message's -- the steady state of decode, where only the
tensors change -- a 1-word "same as before" marker is sent and the receiver reuses its cached
list without unpickling anything.
"""
from __future__ import annotations

import atexit
import mmap
import os
import pickle
import struct
import time
import zlib
from typing import Any

import torch

_APPLIED = False


class ShmObjectChannel:
    """One-directional object channel between two processes of one host: a ring of SLOTS slots
    in a /dev/shm file.  Slot header: seq u64 | ack u64 | len u32 | crc u32 | kind u32.
    Message m (1, 2, ...) uses slot m % SLOTS: the writer waits until that slot's previous
    occupant was acknowledged, writes payload + header fields and, last, seq = m; the reader
    spins on seq == m, checks the CRC (which also covers a weakly ordered host), reads, and sets
    ack = m.  kind: 0 = pickled payload, 1 = same as the previous message, 2 = too large for a
    slot, the payload follows over gloo."""

    SLOTS, SLOT_BYTES, HDR = 16, 16384, 32
    FILE_HDR = 64  # byte 0: receiver attached
    KIND_PICKLE, KIND_SAME, KIND_GLOO = 0, 1, 2

    def __init__(self, name: str, create: bool):
        self.name, self.path = name, os.path.join("/dev/shm", name)
        size = self.FILE_HDR + self.SLOTS * self.SLOT_BYTES
        if create:
            fd = os.open(self.path, os.O_CREAT | os.O_EXCL | os.O_RDWR, 0o600)
            os.ftruncate(fd, size)
        else:
            fd = os.open(self.path, os.O_RDWR)
        try:
            self.mm = mmap.mmap(fd, size)
        finally:
            os.close(fd)
        self.owner = create
        self.unlinked = False
        self.next = 1          # next message index (both sides count on their own)
        self.last: bytes | None = None   # writer: last pickled payload; reader: unused
        self.cached: Any = None          # reader: last unpickled object
        if create:
            atexit.register(self.unlink)
        else:
            self.mm[0] = 1  # tell the writer it may unlink the file: both mappings exist now

    def unlink(self) -> None:
        if self.owner and not self.unlinked:
            self.unlinked = True
            try:
                os.unlink(self.path)
            except OSError:
                pass

    def _off(self, m: int) -> int:
        return self.FILE_HDR + (m % self.SLOTS) * self.SLOT_BYTES

    @staticmethod
    def _spin(cond, what: str, timeout_s: float = 600.0) -> None:
        t0, n = time.monotonic(), 0
        while not cond():
            n += 1
            if n & 0x3ff == 0:
                if time.monotonic() - t0 > timeout_s:
                    raise TimeoutError(f"b200c pp metadata channel: {what} for {timeout_s} s")
                time.sleep(0)  # let other threads of this process run

    def put(self, payload: bytes | None) -> int:
        """payload None = same as before.  Returns the kind that was written."""
        if not self.unlinked and self.mm[0] == 1:
            self.unlink()
        m, off = self.next, self._off(self.next)
        self.next += 1
        # previous occupant of the slot (message m - SLOTS) must have been consumed
        self._spin(lambda: struct.unpack_from("<Q", self.mm, off + 8)[0] ==
                   struct.unpack_from("<Q", self.mm, off)[0], "receiver is 16 messages behind")
        if payload is None:
            kind, body = self.KIND_SAME, b""
        elif len(payload) > self.SLOT_BYTES - self.HDR:
            kind, body = self.KIND_GLOO, b""
        else:
            kind, body = self.KIND_PICKLE, payload
        self.mm[off + self.HDR: off + self.HDR + len(body)] = body
        struct.pack_into("<III", self.mm, off + 16, len(body), zlib.crc32(body), kind)
        struct.pack_into("<Q", self.mm, off, m)
        return kind

    def get(self) -> tuple[int, bytes]:
        m, off = self.next, self._off(self.next)
        self.next += 1
        out: list = []

        def ready() -> bool:
            if struct.unpack_from("<Q", self.mm, off)[0] != m:
                return False
            n, crc, kind = struct.unpack_from("<III", self.mm, off + 16)
            if n > self.SLOT_BYTES - self.HDR:
                return False
            body = bytes(self.mm[off + self.HDR: off + self.HDR + n])
            if zlib.crc32(body) != crc:
                return False
            out[:] = [kind, body]
            return True
        self._spin(ready, "no metadata from the previous stage")
        struct.pack_into("<Q", self.mm, off + 8, m)
        return out[0], out[1]


def _channels(group) -> dict:
    ch = getattr(group, "_b200c_meta_channels", None)
    if ch is None:
        ch = group._b200c_meta_channels = {}
    return ch


def send_meta(group, meta, dst: int) -> str:
    """-> which path carried it: "gloo" (first message of the pair), "shm", "shm-same"."""
    chans = _channels(group)
    payload = pickle.dumps(meta)
    ch = chans.get(("send", dst))
    if ch is None:
        name = f"b200c_pp_{os.getpid()}_{id(group) & 0xffffff:x}_{group.rank_in_group}_{dst}"
        try:
            ch = ShmObjectChannel(name, create=True)
        except OSError:
            ch = None  # no /dev/shm: stay on the reference's gloo path for good
        chans[("send", dst)] = ch or False
        group.send_object({"__b200c_shm__": ch.name if ch else None, "meta": meta}, dst=dst)
        if ch:
            ch.last = payload
        return "gloo"
    if ch is False:
        group.send_object({"__b200c_shm__": None, "meta": meta}, dst=dst)
        return "gloo"
    same = payload == ch.last
    kind = ch.put(None if same else payload)
    ch.last = payload
    if kind == ShmObjectChannel.KIND_GLOO:
        group.send_object(meta, dst=dst)
        return "gloo"
    return "shm-same" if same else "shm"


def recv_meta(group, src: int):
    chans = _channels(group)
    ch = chans.get(("recv", src))
    if ch is None or ch is False:
        first = group.recv_object(src=src)
        if ch is None:
            name = first.get("__b200c_shm__")
            try:
                chans[("recv", src)] = ShmObjectChannel(name, create=False) if name else False
            except OSError:
                # cannot happen for same-node processes that share /dev/shm; refuse to guess
                raise RuntimeError(f"b200c pp metadata channel: cannot open /dev/shm/{name}")
            if chans[("recv", src)]:
                chans[("recv", src)].cached = first["meta"]
        return first["meta"]
    kind, body = ch.get()
    if kind == ShmObjectChannel.KIND_SAME:
        return ch.cached
    meta = group.recv_object(src=src) if kind == ShmObjectChannel.KIND_GLOO else pickle.loads(body)
    ch.cached = meta
    return meta


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
    send_meta(self, meta, dst)
    handles = []
    for (key, _), t in zip([m for m in meta if isinstance(tensor_dict[m[0]], torch.Tensor)], tensors):
        if t.numel() == 0:
            continue
        if _use_slices(self, key, t.numel(), all_gather_group, all_gather_tensors):
            t = t.reshape(ag_size, -1)[ag_rank]
        if t.is_cuda:
            tc = t.contiguous()
            # Both sides pick the transport from what they both know -- the byte count (the
            # receiver derives it from the metadata): mailbox iff it is a multiple of 16.  A
            # source view that is not 16-byte aligned is copied, never re-routed: the peer is
            # already waiting on the mailbox.
            mailbox = (tc.numel() * tc.element_size()) % 16 == 0
            if mailbox and tc.data_ptr() % 16:
                tc = tc.clone()
            if mailbox and hasattr(comm, "isend"):
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
    meta = recv_meta(self, src)
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
        elif hasattr(comm, "irecv") and (shape.numel() * value.dtype.itemsize) % 16 == 0:
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
