"""Center-variable ("parameter server") back-ends for the ADAG / DynSGD trainers.

dist-keras runs a socket parameter server thread in the Spark driver: workers
``commit`` a residual (``center += residual; num_updates += 1`` under a mutex)
and ``pull`` the center.  There is no server process
here; the center variable is replicated on every rank:

* ``CollectiveCenter`` — each communication round is one allreduce-SUM of the
  ranks' residuals, applied to every replica (ADAG *sums* the workers'
  normalised deltas; it does not average them).  Rounds are lock-step, which
  makes the result deterministic; used on CPU/gloo, on one GPU, and wherever
  multicast is unavailable.
* ``MulticastCenter`` (``parallel.peer``) — the center lives in NVSwitch
  multicast memory; a commit is a ``multimem.red.add`` of the scaled residual
  (the switch applies it to every replica) and a pull is a local read, so ranks
  progress without waiting for each other, like the original asynchronous server.

``num_updates`` counts commits over all workers, like
``trainer.parameter_server.num_updates`` (/root/reference/ddl_mnist_aztk.py:224).
"""
from __future__ import annotations

from typing import Optional

import torch

from ..parallel.comm import Communicator, Sum


class CollectiveCenter:
    def __init__(self, comm: Communicator, flat_params: torch.Tensor, staleness_aware: bool = False):
        self.comm = comm
        self.staleness_aware = staleness_aware
        self.center = flat_params.detach().clone()
        self._delta = comm.alloc(flat_params.numel(), torch.float32) if flat_params.is_cuda \
            else torch.zeros_like(flat_params)
        self.local_commits = 0
        self.rounds = 0

    def commit_pull(self, flat_params: torch.Tensor, scale: float, active: bool = True) -> None:
        """``center += sum_r scale_r * (W_r - center_at_last_pull)`` then ``W = center`` on every rank.

        ``scale`` is 1/communication_window for ADAG and 1 for DynSGD; with ``staleness_aware``
        the residual of rank r is additionally divided by (staleness + 1) where, in a lock-step
        round, rank r's commit is the r-th to arrive (staleness = r)."""
        if active:
            s = scale / (self.comm.rank + 1) if self.staleness_aware else scale
            torch.sub(flat_params, self.center, out=self._delta)
            if s != 1.0:
                self._delta.mul_(s)
            self.local_commits += 1
        else:
            self._delta.zero_()
        self.comm.allreduce_(self._delta, Sum)
        self.center.add_(self._delta)
        flat_params.copy_(self.center)
        self.rounds += 1

    def elastic_round(self, flat_params: torch.Tensor, alpha: float, active: bool = True) -> None:
        """Elastic averaging (AEASGD): every worker computes ``E_r = alpha * (W_r - center)``, moves itself towards the
        center (``W_r -= E_r``) and the center towards the workers (``center += sum_r E_r``).  Unlike a commit/pull
        round the worker keeps its own weights."""
        if active:
            torch.sub(flat_params, self.center, out=self._delta)
            self._delta.mul_(alpha)
            flat_params.sub_(self._delta)
            self.local_commits += 1
        else:
            self._delta.zero_()
        self.comm.allreduce_(self._delta, Sum)
        self.center.add_(self._delta)
        self.rounds += 1

    def total_updates(self) -> int:
        t = torch.tensor([self.local_commits], dtype=torch.float64, device=self.center.device)
        if self.comm.size > 1:
            t = t.float()
            self.comm.allreduce_(t, Sum)
        return int(round(float(t.item())))


class MulticastCenter:
    """Asynchronous center variable in NVSwitch multicast memory (K19): a commit is ONE kernel that computes
    ``scale * (W - W_pulled)`` and adds it to every replica with ``multimem.red`` (the switch applies it everywhere),
    then pulls the local replica; ranks never wait for each other, like the original dist-keras server.  Without a
    multicast mapping the same kernel falls back to peer atomics.  ``num_updates`` is a replicated commit counter,
    which is also what DynSGD's staleness term reads."""

    def __init__(self, comm, flat_params: torch.Tensor, staleness_aware: bool = False):
        import ctypes
        from .. import _native as N
        self._N, self._ctypes = N, ctypes
        self.comm = comm
        self.staleness_aware = staleness_aware
        n = flat_params.numel()
        self.n = n
        self.center = comm.alloc(n, torch.float32)
        self.updates = comm.alloc(4, torch.int32)
        self.center_off = comm.offset_of(self.center)
        self.updates_off = comm.offset_of(self.updates)
        self.pulled = flat_params.detach().clone()
        self.last_update = torch.zeros(1, dtype=torch.int32, device=flat_params.device)
        self.center.copy_(flat_params)          # every rank holds identical parameters after the initial broadcast
        torch.cuda.synchronize(flat_params.device)
        comm.barrier()
        torch.cuda.synchronize(flat_params.device)
        self.local_commits = 0
        self.rounds = 0

    def commit_pull(self, flat_params: torch.Tensor, scale: float, active: bool = True) -> None:
        if not active:
            return
        N = self._N
        N.call('ddls_comm_adag_commit', self._ctypes.byref(self.comm.desc), flat_params.data_ptr(), self.pulled.data_ptr(),
               self.center_off, self.updates_off, self.n, scale, int(self.staleness_aware), self.last_update.data_ptr(),
               torch.cuda.current_stream(flat_params.device).cuda_stream)
        torch.autograd.graph.increment_version(flat_params)
        self.local_commits += 1
        self.rounds += 1

    def finalize(self) -> None:
        """Wait until every rank's last commit has landed (call once, after the training loop)."""
        dev = self.center.device
        torch.cuda.synchronize(dev)
        self.comm.barrier()
        torch.cuda.synchronize(dev)

    def total_updates(self) -> int:
        return int(self.updates[0].item())


class ParameterServerView:
    """What ``trainer.parameter_server`` exposes after training."""

    def __init__(self, num_updates: int = 0, center=None):
        self.num_updates = num_updates
        self.center_variable = center

    def get_model(self):
        return self.center_variable
