"""Data-parallel gradient reduction for the spectral weights.

The spectral path shards over data-parallel replicas with no data-path collective; the one exchange per step is the
average of the (complex64) weight gradients, which the reference leaves to DDP with a complex-safe communication hook
(``init_gradient_reduction_hooks``, makani/mpu/mappings.py:104-172: gradients are viewed as real before the
all-reduce).  Here the reduction of a parameter starts the moment its gradient has been accumulated
(``register_post_accumulate_grad_hook``), so on NCCL it overlaps the rest of the backward pass; ``wait()`` joins
them before the optimizer (or the end of the step).
"""
from __future__ import annotations

from typing import Iterable, List

import torch
import torch.distributed as dist


class GradientAverager:
    """Average ``.grad`` of the given parameters over ``group`` as soon as each gradient exists.

        sync = GradientAverager(p for blk in blocks for p in blk.parameters())
        loss.backward(); sync.wait(); optimizer.step()
    """

    def __init__(self, params: Iterable[torch.nn.Parameter], group=None):
        self.group = group
        self.world = dist.get_world_size(group=group) if dist.is_initialized() else 1
        self._pending: List = []
        self._handles = []
        # gloo has no AVG: sum, then scale once in wait()
        self._avg = dist.is_initialized() and dist.get_backend(group) == "nccl"
        self._scaled: List[torch.Tensor] = []
        if self.world > 1:
            for p in params:
                if p.requires_grad:
                    self._handles.append(p.register_post_accumulate_grad_hook(self._hook))

    def _hook(self, p: torch.nn.Parameter):
        g = torch.view_as_real(p.grad) if p.grad.is_complex() else p.grad
        op = dist.ReduceOp.AVG if self._avg else dist.ReduceOp.SUM
        self._pending.append(dist.all_reduce(g, op=op, group=self.group, async_op=True))
        if not self._avg:
            self._scaled.append(g)

    def wait(self):
        for w in self._pending:
            w.wait()
        self._pending.clear()
        for g in self._scaled:
            g.div_(self.world)
        self._scaled.clear()

    def remove(self):
        for h in self._handles:
            h.remove()
        self._handles.clear()
