"""Flat parameter / gradient space.

All parameters of a model are re-pointed at views of ONE contiguous fp32 buffer,
and their ``.grad`` at views of a second one, each segment 16-byte aligned.  The
fused optimizer (``ops.optimizer``), the allreduce kernels and the broadcast all
run over these flat buffers in a single launch; when a peer-memory communicator
is active the buffers live in its symmetric heap so backward writes gradients
straight into collective memory (SURVEY.md §2.5 K6/K12, §5 "Memory").
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Sequence, Tuple

import torch

ALIGN_ELEMS = 4  # 16 B of fp32


class FlatSpace:
    def __init__(self, params: Iterable[torch.nn.Parameter], comm=None, with_grads: bool = True, metric_slots: int = 0):
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError('no trainable parameters')
        dev = self.params[0].device
        for p in self.params:
            if p.dtype != torch.float32:
                raise TypeError('master parameters must be fp32')
            if p.device != dev:
                raise ValueError('all parameters must live on one device')
        self.device = dev
        self.offsets: List[int] = []
        off = 0
        for p in self.params:
            self.offsets.append(off)
            off += (p.numel() + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS
        self.numel = off
        alloc = (lambda n: comm.alloc(n, torch.float32)) if (comm is not None and dev.type == 'cuda') \
            else (lambda n: torch.zeros(n, dtype=torch.float32, device=dev))
        self._alloc = alloc
        # Optional tail of the GRADIENT buffer (K20 of SURVEY.md §2.5): a few extra fp32 slots that ride along with the
        # last gradient bucket through the allreduce (so they come back averaged / summed like the gradients) but are
        # outside the range the optimizer updates and clears.  The parameter buffer has no tail.
        self.metric_slots = (int(metric_slots) + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS
        self.grad_numel = self.numel + self.metric_slots
        self.flat_params = alloc(self.numel)
        self.flat_grads = alloc(self.grad_numel) if with_grads else None
        if with_grads and self.metric_slots:
            self.flat_grads[self.numel:].zero_()
        with torch.no_grad():
            for p, o in zip(self.params, self.offsets):
                view = self.flat_params[o:o + p.numel()].view(p.shape)
                view.copy_(p.data)
                p.data = view
        if with_grads:
            self.attach_grads()

    def new_slot(self) -> torch.Tensor:
        """A zeroed fp32 buffer shaped like the flat parameters, in the same memory (the symmetric heap when a peer
        communicator is active: the sharded data-parallel step addresses optimizer slots by heap offset)."""
        return self._alloc(self.numel)

    def attach_grads(self) -> None:
        for p, o in zip(self.params, self.offsets):
            p.grad = self.flat_grads[o:o + p.numel()].view(p.shape)

    def grads_attached(self) -> bool:
        base = self.flat_grads.data_ptr()
        return all(p.grad is not None and p.grad.data_ptr() == base + 4 * o
                   for p, o in zip(self.params, self.offsets))

    def gather_stray_grads(self) -> None:
        """If autograd replaced a ``.grad`` tensor (e.g. after ``zero_grad(set_to_none=True)``),
        copy it back into the flat buffer and re-attach the view."""
        for p, o in zip(self.params, self.offsets):
            view = self.flat_grads[o:o + p.numel()].view(p.shape)
            if p.grad is None:
                view.zero_()
            elif p.grad.data_ptr() != view.data_ptr():
                view.copy_(p.grad)
            p.grad = view

    def zero_grads(self) -> None:
        self.flat_grads[:self.numel].zero_()

    @property
    def metrics(self) -> Optional[torch.Tensor]:
        """The metric tail of the gradient buffer (None when ``metric_slots == 0``)."""
        return self.flat_grads[self.numel:] if (self.flat_grads is not None and self.metric_slots) else None

    def segments(self) -> List[Tuple[int, int]]:
        return [(o, p.numel()) for p, o in zip(self.params, self.offsets)]
