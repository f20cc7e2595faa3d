"""Data-parallel averaging of the spectral weight gradients over NVLink peer memory - no NCCL kernel.

What it replaces: DDP's all-reduce of the complex64 weight gradients viewed as real (makani/mpu/mappings.py:104-172);
for the named SFNO config that is 8 x 283 MB per step.  An NCCL all-reduce overlapping the backward pass takes SMs and
HBM bandwidth away from the backward kernels (round 1: 12.0 -> 16.2 ms per step at 8 GPUs).  Here the exchange is

    reduce-scatter : one small kernel (sfno_slices_average) forms  (own 1/N slice + the same slice of every peer) / N  in
                     place, reading the peers' slices straight from their CUDA-IPC mapped buffers over NVLink;
    all-gather     : folded into the layout change packed-planar -> complex64 [Cin, Cout, L] that the backward needs
                     anyway (sfno_dhconv_unpack_wgrad_gather): it reads every averaged slice from its owner, so nothing
                     is gathered into a local packed copy and the gradient crosses local HBM as few times as possible
                     (the exchange shares HBM with the backward kernels it overlaps with: at 2 GPUs a variant that
                     pulled the slices with copy engines into staging buffers first - MAKANI_B200_DP_EXCHANGE_COPY=ce,
                     kept for comparison - moved 1.5 GB per block through local HBM and was no faster than NCCL).

All of it is enqueued on a communication stream the moment `dhconv_bwd_w` of a block has been enqueued, ordered between
the ranks by `sfno_peer_barrier` (three per block: gradients complete / averaged slices complete / pulls complete),
and joined by `wait()` before the optimizer.  The packed gradient is written by the GEMM straight into the IPC buffer
(two alternating buffers: the GEMM of block b+2 waits until every peer has finished pulling block b).

Measured (round 2, profiles/r2f_dp_gradient_exchange.txt): 14.70 ms per step at 2 GPUs and 15.97 ms at 4 against 14.38
/ 15.85 ms with the NCCL all-reduce (12.0 ms on one GPU) and bit-identical averaged gradients.  The loss against one GPU
is therefore not NCCL's SM footprint: whichever way the gradients travel, the exchange shares HBM, L2 and NVLink latency
with the backward kernels it overlaps with, and those run 15 - 30 % slower while it is in flight.  bench.py keeps NCCL as
the default and selects this exchange with MAKANI_B200_DP_EXCHANGE=peer.
(Final tree of round 2, 2 GPUs, tools/run_r2_z.sh: 15.3 ms with this exchange, 14.0 ms with NCCL, identical gradients.)

The averaged gradient is assigned to `weight.grad` in `wait()` (added to it if one is already there), not returned
through autograd: its contents only exist once the communication stream has run.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List

import torch
import torch.distributed as dist

from .. import _ext, ops
from . import p2p


class PeerGradientAverager:
    """sync = PeerGradientAverager(blocks); ... loss.backward(); sync.wait(); optimizer.step()"""

    def __init__(self, modules, group=None):
        mods = [m for m in modules if getattr(m, "operator_type", None) == "dhconv" and m.weight.requires_grad]
        if not mods:
            raise ValueError("PeerGradientAverager: no dhconv SpectralConv modules with trainable weights")
        self.group = group if group is not None else dist.group.WORLD
        self.device = mods[0].weight.device
        if not p2p.available(self.group, self.device):
            raise RuntimeError("PeerGradientAverager: peer memory is not available (" + p2p.status()["reason"] + ")")
        self.rank = dist.get_rank(group=self.group)
        self.size = dist.get_world_size(group=self.group)
        if not 2 <= self.size <= 8:
            raise ValueError("PeerGradientAverager: 2..8 ranks on one node")
        self.lib = _ext.load()
        shapes = {tuple(m.weight.shape) for m in mods}
        if len(shapes) != 1:
            raise ValueError("PeerGradientAverager: all spectral weights must have one shape")
        self.Cin, self.Cout, self.L = mods[0].weight.shape
        self.n = ops.wp_elems(self.L, self.Cin, self.Cout)
        per = -(-self.n // self.size)
        self.slice = -(-per // 1024) * 1024                       # floats per slice (16-byte aligned, last one shorter)
        self.bounds = [(min(r * self.slice, self.n), min((r + 1) * self.slice, self.n)) for r in range(self.size)]
        self.buffers = p2p.PeerBuffers(self.n, self.group, self.device)       # two alternating gradient buffers, IPC mapped
        self.fence = p2p.PeerFence(self.group, self.device)                   # own flags: lives on the communication stream
        self.staging = torch.empty((self.size - 1) * self.slice, dtype=torch.float32, device=self.device)
        self.stream = torch.cuda.Stream(device=self.device)
        self.use_ce = os.environ.get("MAKANI_B200_DP_EXCHANGE_COPY", "direct") == "ce"
        self.free = [None, None]                                  # event: buffer b may be overwritten by the next GEMM
        self.count = 0
        self._ready: List = []
        self.modules = mods
        for m in mods:
            m._grad_exchange = self

    # ---- called from SpectralConv's backward (main stream)
    def gwp_out(self) -> torch.Tensor:
        b = self.count & 1
        if self.free[b] is not None:
            torch.cuda.current_stream().wait_event(self.free[b])
        return self.buffers.local[b]

    def reduce(self, mod, gwp: torch.Tensor):
        b = self.count & 1
        self.count += 1
        assert gwp.data_ptr() == self.buffers.local[b].data_ptr()
        main = torch.cuda.current_stream()
        produced = torch.cuda.Event()
        produced.record(main)
        lib, n4 = self.lib, 4
        me, size = self.rank, self.size
        lo, hi = self.bounds[me]
        with torch.cuda.stream(self.stream):
            cs = C.c_void_p(self.stream.cuda_stream)
            self.stream.wait_event(produced)
            self.fence()                                           # 1: every rank's packed gradient of this block is complete
            srcs, k = [], 0
            for p in range(size):
                if p == me or hi <= lo:
                    continue
                src = self.buffers.peer_ptrs[b][p] + lo * n4
                if self.use_ce:                                    # pull into local staging with a copy engine first
                    dst = self.staging.data_ptr() + k * self.slice * n4
                    _ext.check(lib.sfno_peer_copy(C.c_void_p(dst), C.c_void_p(src), (hi - lo) * n4, cs), "sfno_peer_copy")
                    src = dst
                srcs.append(src)
                k += 1
            if hi > lo:
                arr = (C.c_void_p * max(1, len(srcs)))(*srcs)
                _ext.check(lib.sfno_slices_average(C.c_void_p(gwp.data_ptr() + lo * n4), C.cast(arr, C.c_void_p), len(srcs),
                                                   hi - lo, 1.0 / size, cs), "sfno_slices_average")
            self.fence()                                           # 2: every rank's averaged slice is complete
            gw = torch.empty((self.Cin, self.Cout, self.L), dtype=torch.complex64, device=self.device)
            if self.use_ce:
                for p in range(size):
                    plo, phi = self.bounds[p]
                    if p == me or phi <= plo:
                        continue
                    _ext.check(lib.sfno_peer_copy(C.c_void_p(gwp.data_ptr() + plo * n4),
                                                  C.c_void_p(self.buffers.peer_ptrs[b][p] + plo * n4), (phi - plo) * n4, cs),
                               "sfno_peer_copy")
                _ext.check(lib.sfno_dhconv_unpack_wgrad(C.c_void_p(gwp.data_ptr()), self.Cin, self.Cout, self.L,
                                                        C.c_void_p(gw.data_ptr()), cs), "sfno_dhconv_unpack_wgrad")
            else:                                                  # slice p is read from rank p's buffer (own slice locally)
                bases = (C.c_void_p * size)(*[self.buffers.peer_ptrs[b][p] for p in range(size)])
                _ext.check(lib.sfno_dhconv_unpack_wgrad_gather(C.cast(bases, C.c_void_p), size, self.slice, self.Cin,
                                                               self.Cout, self.L, C.c_void_p(gw.data_ptr()), cs),
                           "sfno_dhconv_unpack_wgrad_gather")
            self.fence()                                           # 3: nobody reads this buffer any more
            self.free[b] = torch.cuda.Event()
            self.free[b].record(self.stream)
        self._ready.append((mod, gw))

    # ---- before the optimizer / at the end of the step
    def wait(self):
        main = torch.cuda.current_stream()
        main.wait_stream(self.stream)
        for mod, gw in self._ready:
            gw.record_stream(main)
            if mod.weight.grad is None:
                mod.weight.grad = gw
            else:
                mod.weight.grad.add_(gw)
        self._ready.clear()

    def remove(self):
        self.wait()
        for m in self.modules:
            if getattr(m, "_grad_exchange", None) is self:
                m._grad_exchange = None
