"""Communicators: the collective backend behind ``allreduce/broadcast/allgather/barrier``.

The reference moves parameters as pickles over Python TCP sockets between
dist-keras workers and a parameter-server thread.  Here:

* ``PeerMemoryComm`` (GPUs, world > 1): hand-written sm_100a kernels over a VMM
  symmetric heap mapped into every rank (``parallel.peer``): one-shot push
  allreduce, two-shot reduce-scatter/all-gather, NVLS ``multimem`` allreduce and
  a multicast-store broadcast, each fused with the 1/N (or 1/window) scale.
* ``TorchDistComm``: ``torch.distributed`` (gloo on CPU; NCCL as the labelled
  baseline on GPU, ``DDLS_COMM=nccl``).
* ``LocalComm``: world of one.
"""
from __future__ import annotations

from enum import Enum
from typing import List, Optional

import torch
import torch.distributed as dist


class ReduceOp(Enum):
    Average = 'average'
    Sum = 'sum'


Average = ReduceOp.Average
Sum = ReduceOp.Sum


class Communicator:
    name = 'base'

    def __init__(self, rank: int, size: int, device: torch.device):
        self.rank, self.size, self.device = rank, size, device

    # memory ------------------------------------------------------------
    def alloc(self, numel: int, dtype: torch.dtype = torch.float32) -> torch.Tensor:
        """Buffer usable as a collective operand (symmetric memory when the backend has a heap)."""
        return torch.zeros(numel, dtype=dtype, device=self.device)

    # collectives -------------------------------------------------------
    def allreduce_(self, t: torch.Tensor, op: ReduceOp = Average, prescale: float = 1.0) -> torch.Tensor:
        raise NotImplementedError

    def broadcast_(self, t: torch.Tensor, root: int = 0) -> torch.Tensor:
        raise NotImplementedError

    def allgather(self, t: torch.Tensor) -> torch.Tensor:
        raise NotImplementedError

    def barrier(self) -> None:
        raise NotImplementedError

    def allgather_object(self, obj) -> List:
        raise NotImplementedError

    def broadcast_object(self, obj, root: int = 0):
        raise NotImplementedError

    def close(self) -> None:
        pass


class LocalComm(Communicator):
    name = 'local'

    def allreduce_(self, t, op=Average, prescale=1.0):
        if prescale != 1.0:
            t.mul_(prescale)
        return t

    def broadcast_(self, t, root=0):
        return t

    def allgather(self, t):
        return t.unsqueeze(0).clone()

    def barrier(self):
        return None

    def allgather_object(self, obj):
        return [obj]

    def broadcast_object(self, obj, root=0):
        return obj


class TorchDistComm(Communicator):
    """``torch.distributed`` collectives: gloo (CPU plumbing tests; CUDA tensors bounce through the host) or NCCL (the
    baseline a product path must beat, ``DDLS_COMM=nccl``)."""

    name = 'torch.distributed'

    def __init__(self, rank, size, device, group=None):
        super().__init__(rank, size, device)
        self.group = group
        self._host_bounce = device.type == 'cuda' and dist.get_backend(group) != 'nccl'

    def _host(self, t):
        return t.detach().cpu() if self._host_bounce else t

    def allreduce_(self, t, op=Average, prescale=1.0):
        if prescale != 1.0:
            t.mul_(prescale)
        h = self._host(t)
        dist.all_reduce(h, op=dist.ReduceOp.SUM, group=self.group)
        if h is not t:
            t.copy_(h)
        if op is Average:
            t.div_(self.size)
        return t

    def broadcast_(self, t, root=0):
        h = self._host(t)
        dist.broadcast(h, src=root, group=self.group)
        if h is not t:
            t.copy_(h)
        return t

    def allgather(self, t):
        h = self._host(t.contiguous())
        out = [torch.empty_like(h) for _ in range(self.size)]
        dist.all_gather(out, h, group=self.group)
        return torch.stack(out).to(t.device)

    def barrier(self):
        if self.device.type == 'cuda':
            torch.cuda.synchronize(self.device)
        if dist.get_backend(self.group) == 'nccl':
            dist.barrier(group=self.group, device_ids=[self.device.index])
        else:
            dist.barrier(group=self.group)

    def allgather_object(self, obj):
        out = [None] * self.size
        dist.all_gather_object(out, obj, group=self.group)
        return out

    def broadcast_object(self, obj, root=0):
        box = [obj]
        dist.broadcast_object_list(box, src=root, group=self.group)
        return box[0]


def make_communicator(state, prefer: str = 'auto') -> Communicator:
    if state.size == 1:
        return LocalComm(0, 1, state.device)
    if state.device.type == 'cuda' and prefer in ('auto', 'peer'):
        try:
            from .peer import PeerMemoryComm
            return PeerMemoryComm(state.rank, state.size, state.device, state.local_rank, state.local_size)
        except Exception as e:  # noqa: BLE001
            if prefer == 'peer':
                raise
            from ..utils.logging import get_logger
            get_logger().warning('peer-memory communicator unavailable (%s); using torch.distributed', e)
    group = None
    if state.device.type == 'cuda' and dist.get_backend() != 'nccl':
        try:                                    # a GPU fallback / baseline communicator wants NCCL, not host bounces
            group = dist.new_group(backend='nccl')
        except Exception:  # noqa: BLE001
            group = None
    return TorchDistComm(state.rank, state.size, state.device, group=group)
