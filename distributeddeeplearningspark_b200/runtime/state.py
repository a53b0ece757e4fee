"""Process-global runtime: ``init/rank/size/local_rank/local_size/shutdown`` and topology.

Horovod-shaped surface required by BASELINE.json (the reference itself gets its
worker count from ``num_executors * num_processes``,
/root/reference/ddl_mnist_aztk.py:49-53).  ``torch.distributed`` with the **gloo**
backend (a TCP store + CPU collectives) is used for *plumbing only* — rendezvous,
handle exchange, host barriers, object gathers — on CPUs and GPUs alike, so the
product path never initialises NCCL.  The gradient/parameter collectives on GPUs
run through the hand-written peer-memory kernels in ``parallel.comm``; NCCL is
only created when asked for by name (``DDLS_COMM=nccl``, the labelled baseline,
and the comparison arms of the benchmarks).
"""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass, field
from typing import List, Optional

import torch
import torch.distributed as dist

from ..utils.logging import get_logger


@dataclass
class Topology:
    device_count: int = 0
    device_name: str = 'cpu'
    sm_count: int = 0
    p2p_matrix: List[List[bool]] = field(default_factory=list)
    multicast_supported: bool = False
    nvlink_all_to_all: bool = False


@dataclass
class _State:
    initialized: bool = False
    rank: int = 0
    size: int = 1
    local_rank: int = 0
    local_size: int = 1
    device: torch.device = torch.device('cpu')
    backend: str = 'none'
    owns_pg: bool = False
    topology: Optional[Topology] = None
    comm: object = None  # parallel.comm.Communicator


_S = _State()


def _probe_topology(device: torch.device) -> Topology:
    if device.type != 'cuda':
        return Topology()
    n = torch.cuda.device_count()
    props = torch.cuda.get_device_properties(device)
    mat = [[(i == j) or bool(torch.cuda.can_device_access_peer(i, j)) for j in range(n)] for i in range(n)]
    mc = False
    try:
        from .. import _native
        mc = bool(_native.lib().ddls_multicast_supported(device.index or 0))
    except Exception:
        mc = False
    return Topology(device_count=n, device_name=props.name, sm_count=props.multi_processor_count,
                    p2p_matrix=mat, multicast_supported=mc,
                    nvlink_all_to_all=all(all(r) for r in mat) and n > 1)


def init(backend: Optional[str] = None, timeout_s: float = 300.0, comm: Optional[str] = None) -> None:
    """Join the job.  Reads ``RANK/WORLD_SIZE/LOCAL_RANK/MASTER_ADDR/MASTER_PORT`` (set by
    ``spark.run`` or ``torchrun``); a lone process becomes rank 0 of 1."""
    if _S.initialized:
        return
    env = os.environ
    _S.rank = int(env.get('RANK', env.get('DDLS_RANK', '0')))
    _S.size = int(env.get('WORLD_SIZE', env.get('DDLS_SIZE', '1')))
    _S.local_rank = int(env.get('LOCAL_RANK', env.get('DDLS_LOCAL_RANK', str(_S.rank))))
    _S.local_size = int(env.get('LOCAL_WORLD_SIZE', env.get('DDLS_LOCAL_SIZE', str(_S.size))))
    use_cuda = torch.cuda.is_available() and env.get('DDLS_FORCE_CPU', '0') != '1'
    if use_cuda:
        idx = _S.local_rank % torch.cuda.device_count()
        torch.cuda.set_device(idx)
        _S.device = torch.device('cuda', idx)
    else:
        _S.device = torch.device('cpu')
    prefer = comm or env.get('DDLS_COMM', 'auto')
    _S.backend = backend or env.get('DDLS_PG_BACKEND') or ('nccl' if (use_cuda and prefer == 'nccl') else 'gloo')
    if _S.size > 1:
        if dist.is_initialized():
            _S.owns_pg = False
        else:
            env.setdefault('MASTER_ADDR', '127.0.0.1')
            env.setdefault('MASTER_PORT', '29511')
            kw = {}
            if use_cuda and _S.backend == 'nccl':
                kw['device_id'] = _S.device
            dist.init_process_group(backend=_S.backend, rank=_S.rank, world_size=_S.size,
                                    timeout=datetime.timedelta(seconds=timeout_s), **kw)
            _S.owns_pg = True
    _S.topology = _probe_topology(_S.device)
    _S.initialized = True
    from ..parallel.comm import make_communicator
    _S.comm = make_communicator(_S, prefer=prefer)
    log = get_logger()
    if _S.rank == 0:
        t = _S.topology
        log.info('init: size=%d backend=%s device=%s (%s, %d SMs) comm=%s p2p_all=%s multicast=%s', _S.size,
                 _S.backend, _S.device, t.device_name, t.sm_count, type(_S.comm).__name__,
                 t.nvlink_all_to_all, t.multicast_supported)


def shutdown() -> None:
    if not _S.initialized:
        return
    try:
        if _S.comm is not None:
            _S.comm.close()
    finally:
        _S.comm = None
        if _S.owns_pg and dist.is_initialized():
            try:
                dist.destroy_process_group()
            except Exception:
                pass
        _S.initialized = False
        _S.owns_pg = False
        _S.rank, _S.size, _S.local_rank, _S.local_size = 0, 1, 0, 1


def is_initialized() -> bool:
    return _S.initialized


def _need():
    if not _S.initialized:
        raise RuntimeError('call init() first')
    return _S


def rank() -> int:
    return _need().rank


def size() -> int:
    return _need().size


def local_rank() -> int:
    return _need().local_rank


def local_size() -> int:
    return _need().local_size


def device() -> torch.device:
    return _need().device


def topology() -> Topology:
    return _need().topology


def communicator():
    return _need().comm
