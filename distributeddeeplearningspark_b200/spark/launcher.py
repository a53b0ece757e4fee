"""``spark.run(fn, args, kwargs, num_proc)`` — Horovod-on-Spark-shaped launcher.

One task per GPU inside a barrier stage.  Each task learns its rank from the
partition id, finds its local rank from the task addresses, pins
``cuda:<local_rank>``, agrees on a rendezvous endpoint through the stage's
``allGather`` (rank 0 proposes host:port), exports the usual
``RANK/WORLD_SIZE/LOCAL_RANK/MASTER_ADDR/MASTER_PORT`` variables and calls
``fn(*args, **kwargs)``.  The per-rank return values come back ordered by rank,
like ``horovod.spark.run``.  In the reference the equivalent fan-out is
dist-keras' ``rdd.mapPartitionsWithIndex(worker.train).collect()``
(/root/reference/ddl_mnist_aztk.py:216-219).
"""
from __future__ import annotations

import os
from typing import Any, Callable, Dict, List, Optional, Sequence

from ..utils.misc import determine_host_address, free_port
from .backends import InProcessBackend, LocalBarrierBackend, PySparkBarrierBackend, pyspark_available


def _select_backend(backend, num_proc: Optional[int]):
    if backend is None or backend == 'auto':
        backend = os.environ.get('DDLS_SPARK_BACKEND', 'local')
    if isinstance(backend, str):
        if backend == 'local':
            return LocalBarrierBackend()
        if backend == 'inprocess':
            return InProcessBackend()
        if backend == 'pyspark':
            return PySparkBarrierBackend()
        raise ValueError(f'unknown backend {backend!r}')
    return backend


def _make_task(fn: Callable, args: Sequence, kwargs: Dict, env: Dict[str, str], use_gpu: Optional[bool]):
    def task(ctx):
        rank, size = ctx.partition_id(), ctx.num_tasks()
        addrs = ctx.task_addresses()
        my_addr = addrs[rank]
        local_peers = [i for i, a in enumerate(addrs) if a == my_addr]
        local_rank, local_size = local_peers.index(rank), len(local_peers)
        proposal = ''
        if rank == 0:
            host = '127.0.0.1' if all(a == my_addr for a in addrs) else determine_host_address()
            proposal = f'{host}:{free_port(host if host == "127.0.0.1" else "")}'
        master = ctx.all_gather(proposal)[0]
        host, port = master.rsplit(':', 1)
        os.environ.update(env)
        os.environ.update({'RANK': str(rank), 'WORLD_SIZE': str(size), 'LOCAL_RANK': str(local_rank),
                           'LOCAL_WORLD_SIZE': str(local_size), 'MASTER_ADDR': host, 'MASTER_PORT': port,
                           'DDLS_RANK': str(rank), 'DDLS_SIZE': str(size),
                           'DDLS_LOCAL_RANK': str(local_rank), 'DDLS_LOCAL_SIZE': str(local_size)})
        import torch
        gpu = torch.cuda.is_available() if use_gpu is None else use_gpu
        if gpu:
            torch.cuda.set_device(local_rank % max(1, torch.cuda.device_count()))
        from .. import runtime
        was_init = runtime.is_initialized()
        try:
            return fn(*args, **(kwargs or {}))
        finally:
            if runtime.is_initialized() and not was_init:
                runtime.shutdown()
    return task


def run(fn: Callable, args: Sequence = (), kwargs: Optional[Dict] = None, num_proc: Optional[int] = None,
        backend=None, env: Optional[Dict[str, str]] = None, use_gpu: Optional[bool] = None,
        verbose: int = 0) -> List[Any]:
    """Run ``fn`` on ``num_proc`` ranks; returns ``[fn result of rank 0, rank 1, ...]``."""
    be = _select_backend(backend, num_proc)
    n = int(num_proc) if num_proc is not None else int(be.default_parallelism())
    if n < 1:
        raise ValueError('num_proc must be >= 1')
    if verbose:
        print(f'[ddls.spark.run] launching {n} rank(s) on {type(be).__name__}')
    task = _make_task(fn, tuple(args), dict(kwargs or {}), dict(env or {}), use_gpu)
    return be.run_barrier_stage(n, task)
