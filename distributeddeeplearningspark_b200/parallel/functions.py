"""Collective functions of the Horovod-shaped API: ``allreduce / allgather / broadcast / barrier``,
``broadcast_parameters`` and ``broadcast_optimizer_state``.

dist-keras hands every worker the initial weights inside the pickled task closure plus a first
``pull()`` from the parameter server; the equivalent here is a rank-0 broadcast
of the flat parameter buffer (one multicast-store kernel on GPUs)."""
from __future__ import annotations

from typing import Dict, Iterable, Optional, Union

import torch

from .. import runtime
from ..optim.keras_optimizers import _FlatOptimizer
from .comm import Average, ReduceOp, Sum


def _comm():
    if not runtime.is_initialized():
        runtime.init()
    return runtime.communicator()


def allreduce(tensor: torch.Tensor, op: ReduceOp = Average, name: Optional[str] = None) -> torch.Tensor:
    out = tensor.detach().clone().contiguous()
    _comm().allreduce_(out.view(-1), op)
    return out


def allreduce_(tensor: torch.Tensor, op: ReduceOp = Average, name: Optional[str] = None) -> torch.Tensor:
    if not tensor.is_contiguous():
        raise ValueError('allreduce_ needs a contiguous tensor')
    _comm().allreduce_(tensor.view(-1), op)
    return tensor


def allgather(tensor: torch.Tensor, name: Optional[str] = None) -> torch.Tensor:
    """Concatenate every rank's tensor along dim 0 (equal shapes)."""
    g = _comm().allgather(tensor.contiguous())
    return g.reshape((-1,) + tuple(tensor.shape[1:])) if tensor.dim() > 0 else g.reshape(-1)


def broadcast(tensor: torch.Tensor, root_rank: int = 0, name: Optional[str] = None) -> torch.Tensor:
    out = tensor.detach().clone().contiguous()
    _comm().broadcast_(out, root_rank)
    return out


def broadcast_(tensor: torch.Tensor, root_rank: int = 0, name: Optional[str] = None) -> torch.Tensor:
    _comm().broadcast_(tensor, root_rank)
    return tensor


def barrier() -> None:
    _comm().barrier()


def allgather_object(obj):
    return _comm().allgather_object(obj)


def broadcast_object(obj, root_rank: int = 0):
    return _comm().broadcast_object(obj, root_rank)


def broadcast_parameters(params: Union[Dict[str, torch.Tensor], Iterable, torch.nn.Module, _FlatOptimizer],
                         root_rank: int = 0) -> None:
    """Make every rank's parameters bit-identical to ``root_rank``'s.  Accepts a module, a
    ``state_dict()``, ``named_parameters()`` or a flat optimizer (single launch over its flat buffer)."""
    comm = _comm()
    if isinstance(params, _FlatOptimizer):
        comm.broadcast_(params.flat.flat_params, root_rank)
        return
    if isinstance(params, torch.nn.Module):
        params = params.state_dict()
    items = sorted(params.items()) if isinstance(params, dict) else list(params)
    seen = set()
    for _, t in items:
        if not torch.is_tensor(t):
            continue
        base = t.data if isinstance(t, torch.nn.Parameter) else t
        key = (base.untyped_storage().data_ptr(), base.storage_offset(), base.numel())
        if key in seen:
            continue
        seen.add(key)
        if base.is_contiguous():
            comm.broadcast_(base, root_rank)
        else:
            tmp = base.contiguous()
            comm.broadcast_(tmp, root_rank)
            base.copy_(tmp)


def broadcast_optimizer_state(optimizer, root_rank: int = 0) -> None:
    comm = _comm()
    inner = getattr(optimizer, 'inner', optimizer)
    if isinstance(inner, _FlatOptimizer):
        # The root's slots are taken as complete and authoritative (a freshly constructed optimizer, or one that has just
        # loaded a checkpoint).  Mid-training, when the sharded data-parallel step keeps each slice of m / v up to date
        # only on its owner, call ``optimizer.gather_slots()`` on every rank first.
        comm.broadcast_(inner.step_count, root_rank)
        for t in inner.slots().values():
            comm.broadcast_(t, root_rank)
        return
    sd = comm.broadcast_object(inner.state_dict() if comm.rank == root_rank else None, root_rank)
    if comm.rank != root_rank:
        inner.load_state_dict(sd)
