"""Peer-memory exchange for the h-sharded (polar) distributed SHT.

The reference transposes between its two data distributions with ``split`` + ``contiguous`` per chunk,
``dist.all_to_all`` and ``cat`` (thd.distributed_transpose_polar; call sites makani/mpu/layers.py:73-100), i.e. three
passes over the data plus the wire.  On one NVLink / NVSwitch node every GPU can store straight into every other
GPU's memory, so here the transposition is ONE kernel (``sfno_rows_scatter``): it reads the local internal-layout
tensor (XM or SP) once and writes each row to its final position in the owner's receive buffer - a local buffer for
the own part, CUDA-IPC mapped peer buffers for the rest.  No packing, no unpacking, no complex64 round trip.

Synchronisation is stream ordered and needs no host sync:
   write kernel (all ranks)  ->  peer barrier on the same stream  ->  consumers of the receive buffer
The barrier (``sfno_peer_barrier``: every rank stores an epoch into every rank's flag array in peer memory with
system-scope release and spins on its own array with acquire loads) cannot complete on rank r before every rank has
reached its own, i.e. before every rank's write kernel has finished, so the buffer is complete when the consumer
starts.  ``MAKANI_B200_P2P_FENCE=nccl`` uses a one-element NCCL all-reduce instead.  Receive buffers are used
alternately (two per layout); a buffer is written again only two exchanges later, after an intervening barrier that
every rank enters behind its own consumer of the earlier contents - so no second barrier is needed for the
write-after-read hazard.

CUDA graphs (the reference captures forward + backward, makani/utils/trainer.py:84-152): the barrier number lives in
device memory (the kernel increments it), so a captured barrier replays correctly; the buffer alternation, however,
is decided on the host and is baked into a captured graph - a graph with an odd number of exchanges per layout would
write the same buffer at its end and again at the start of the next replay with no barrier in between.  While a stream
is capturing, every exchange is therefore preceded by one more barrier (``pre_write``): a rank enters it behind its own
last consumer of the buffer, so nobody overwrites contents that are still being read.
"""
from __future__ import annotations

import ctypes as C
import os
import socket
from typing import Dict, List, Tuple

import torch
import torch.distributed as dist

from .. import _ext


class _DevPtr:
    """Expose a raw device pointer to torch (``torch.as_tensor``) through the CUDA array interface."""

    def __init__(self, ptr: int, nfloats: int):
        self.__cuda_array_interface__ = {"shape": (nfloats,), "typestr": "<f4", "data": (ptr, False), "version": 2}


class PeerBuffers:
    """Two (alternating) zero-initialised float buffers per rank of ``group``, each mapped into every rank."""

    def __init__(self, nfloats: int, group, device: torch.device):
        lib = _ext.load()
        self.group = group
        self.rank = dist.get_rank(group=group)
        self.size = dist.get_world_size(group=group)
        self.nfloats = nfloats
        self.local_ptrs: List[int] = []
        self.local: List[torch.Tensor] = []
        self.peer_ptrs: List[List[int]] = []          # [buffer][rank] -> device pointer valid in THIS process
        self._use = 0
        handles = []
        with torch.cuda.device(device):
            for _ in range(2):
                p = C.c_void_p()
                h = C.create_string_buffer(64)
                _ext.check(lib.sfno_ipc_alloc(nfloats * 4, C.byref(p), h), "sfno_ipc_alloc")
                self.local_ptrs.append(p.value)
                self.local.append(torch.as_tensor(_DevPtr(p.value, nfloats), device=device))
                handles.append(bytes(h.raw))
            gathered = [None] * self.size
            dist.all_gather_object(gathered, handles, group=group)
            for b in range(2):
                ptrs = []
                for r in range(self.size):
                    if r == self.rank:
                        ptrs.append(self.local_ptrs[b])
                        continue
                    q = C.c_void_p()
                    _ext.check(lib.sfno_ipc_open(gathered[r][b], C.byref(q)), "sfno_ipc_open")
                    ptrs.append(q.value)
                self.peer_ptrs.append(ptrs)

    def next(self) -> int:
        b = self._use & 1
        self._use += 1
        return b


class PeerFence:
    """Stream-ordered barrier of the ranks of ``group`` (see module docstring)."""

    def __init__(self, group, device: torch.device):
        lib = _ext.load()
        self.group = group
        self.rank = dist.get_rank(group=group)
        self.size = dist.get_world_size(group=group)
        self.use_nccl = os.environ.get("MAKANI_B200_P2P_FENCE", "peer") == "nccl"
        self.barriers = 0                      # issued from this process (statistics only; the number itself is on the device)
        self._flag = torch.zeros(1, dtype=torch.float32, device=device)
        with torch.cuda.device(device):
            p = C.c_void_p()
            h = C.create_string_buffer(64)
            _ext.check(lib.sfno_ipc_alloc(256, C.byref(p), h), "sfno_ipc_alloc")
            gathered = [None] * self.size
            dist.all_gather_object(gathered, bytes(h.raw), group=group)
            ptrs = []
            for r in range(self.size):
                if r == self.rank:
                    ptrs.append(p.value)
                    continue
                q = C.c_void_p()
                _ext.check(lib.sfno_ipc_open(gathered[r], C.byref(q)), "sfno_ipc_open")
                ptrs.append(q.value)
        self._ptrs = (C.c_void_p * self.size)(*ptrs)

    def __call__(self):
        """Everything all ranks enqueued before this point is complete and visible to what follows (stream ordered)."""
        if self.use_nccl:
            dist.all_reduce(self._flag, group=self.group)
            return
        self.barriers += 1
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        _ext.check(_ext.load().sfno_peer_barrier(C.cast(self._ptrs, C.c_void_p), self.rank, self.size, 0, stream),
                   "sfno_peer_barrier")

    def pre_write(self):
        """Called before a kernel writes into receive buffers: under CUDA-graph capture the host-side buffer alternation
        cannot be relied on (module docstring), so order the write behind every rank's last consumer."""
        if torch.cuda.is_current_stream_capturing():
            self()


_FENCES: Dict[Tuple, PeerFence] = {}


def fence(group, device: torch.device) -> PeerFence:
    k = (id(group), device.index)
    f = _FENCES.get(k)
    if f is None:
        f = PeerFence(group, device)
        _FENCES[k] = f
    return f


_POOL: Dict[Tuple, PeerBuffers] = {}
_STATUS = {"enabled": None, "reason": "not initialised", "exchanges": 0}


def status() -> dict:
    """{'enabled': bool | None, 'reason': str, 'exchanges': int} - which path the distributed transforms take."""
    return dict(_STATUS)


def available(group, device: torch.device) -> bool:
    """Peer-memory path usable for ``group``: CUDA tensors, NCCL backend, all ranks on one host, not disabled."""
    if _STATUS["enabled"] is not None:
        return _STATUS["enabled"]
    ok, why = True, "peer memory (CUDA IPC over NVLink)"
    if os.environ.get("MAKANI_B200_P2P", "1") == "0":
        ok, why = False, "disabled by MAKANI_B200_P2P=0"
    elif device.type != "cuda":
        ok, why = False, "not a CUDA device"
    elif dist.get_backend(group) != "nccl":
        ok, why = False, f"backend {dist.get_backend(group)}"
    else:
        hosts = [None] * dist.get_world_size(group=group)
        dist.all_gather_object(hosts, socket.gethostname(), group=group)
        if len(set(hosts)) != 1:
            ok, why = False, "ranks span several hosts"
    _STATUS["enabled"], _STATUS["reason"] = ok, why
    return ok


def buffers(key: Tuple, nfloats: int, group, device: torch.device) -> PeerBuffers:
    """Receive buffers for one destination layout (``key`` names the layout: pad entries are zero at allocation and
    never written, which the consumers rely on, so buffers are not shared between layouts)."""
    k = (key, nfloats, id(group), device.index)
    pb = _POOL.get(k)
    if pb is None:
        pb = PeerBuffers(nfloats, group, device)
        _POOL[k] = pb
    return pb


def desc_array(descs: List[_ext.CopyDesc]):
    return (_ext.CopyDesc * len(descs))(*descs)


# MAKANI_B200_P2P_COPY=ce moves the exchanges whose rows form a plain strided box (the two XM exchanges: latitude <->
# channel) with the copy engines (sfno_rows_copy_ce) instead of the SM kernel; default "sm" (measured)
_COPY = os.environ.get("MAKANI_B200_P2P_COPY", "sm")


def scatter(src: torch.Tensor, arr, boxed: bool = False):
    """arr: ctypes array of sfno_copy_desc (desc_array); one kernel moves every row to its owner.  boxed: the
    descriptors are plain strided boxes (no triangular skip), eligible for the copy-engine variant."""
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if boxed and _COPY == "ce":
        _ext.check(_ext.load().sfno_rows_copy_ce(C.c_void_p(src.data_ptr()), C.cast(arr, C.c_void_p), len(arr), stream),
                   "sfno_rows_copy_ce")
    else:
        _ext.check(_ext.load().sfno_rows_scatter(C.c_void_p(src.data_ptr()), C.cast(arr, C.c_void_p), len(arr), stream),
                   "sfno_rows_scatter")
    _STATUS["exchanges"] += 1
