"""Per-rank training step engine: forward + loss + backward (+ allreduce) + optimizer.

This is what dist-keras' workers do with ``model.train_on_batch(X, Y)``
(SURVEY.md §3.2 hot loop).  On CUDA the whole step — fused forward/backward
kernels, the bucketed allreduce and the fused optimizer kernel — is captured
once in a CUDA graph and replayed per minibatch, because these models are
launch-latency bound.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

import torch

from ..models.sequential import Sequential
from ..optim import get_optimizer
from ..utils.misc import env_flag

EPS = 1e-7


def categorical_crossentropy(probs: torch.Tensor, onehot: torch.Tensor) -> torch.Tensor:
    """Keras form on softmax outputs (renormalise, clip, -sum y log p), mean over the batch."""
    p = probs / probs.sum(dim=-1, keepdim=True)
    p = p.clamp(EPS, 1.0 - EPS)
    return -(onehot * p.log()).sum(dim=-1).mean()


def mean_squared_error(pred: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    return ((pred - target.reshape(pred.shape)) ** 2).mean(dim=-1).mean()


LOSSES: Dict[str, Callable] = {
    'categorical_crossentropy': categorical_crossentropy,
    'mean_squared_error': mean_squared_error,
    'mse': mean_squared_error,
}


def get_loss(name: str) -> Callable:
    if callable(name):
        return name
    if name not in LOSSES:
        raise ValueError(f'unsupported loss {name!r}; have {sorted(LOSSES)}')
    return LOSSES[name]


class StepEngine:
    """Owns model + optimizer on one device and runs training steps.

    ``optimizer`` may be a name (Keras string), a flat optimizer, or a
    ``DistributedOptimizer`` wrapping one.  ``step(x, y)`` returns the loss as a
    device tensor without synchronising."""

    def __init__(self, model: Sequential, optimizer='adam', loss: str = 'mean_squared_error', device=None,
                 comm=None, distributed: bool = False, use_cuda_graph: Optional[bool] = None,
                 optimizer_kwargs: Optional[dict] = None, dist_kwargs: Optional[dict] = None):
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        self.model = model.to(self.device)
        self.loss_name = loss if isinstance(loss, str) else getattr(loss, '__name__', 'custom')
        self.loss_fn = get_loss(loss)
        base = get_optimizer(optimizer, self.model.parameters(), comm=comm, **(optimizer_kwargs or {})) \
            if isinstance(optimizer, str) else optimizer
        if distributed and not hasattr(base, 'synchronize'):
            from ..parallel import DistributedOptimizer
            base = DistributedOptimizer(base, **(dist_kwargs or {}))
        self.optimizer = base
        self.cuda = self.device.type == 'cuda'
        if use_cuda_graph is None:
            use_cuda_graph = self.cuda and env_flag('DDLS_CUDA_GRAPH', True)
        self.use_cuda_graph = bool(use_cuda_graph) and self.cuda
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._static_x: Optional[torch.Tensor] = None
        self._static_y: Optional[torch.Tensor] = None
        self._static_loss: Optional[torch.Tensor] = None
        self._warm = 0
        self.steps = 0
        inner = getattr(self.optimizer, 'inner', self.optimizer)
        self._inner = inner
        self._direct = None          # autograd-free fused step (ops.fused.DirectStep), built lazily
        self._direct_checked = False
        self.use_direct = self.cuda and env_flag('DDLS_DIRECT_STEP', True)
        # the fused optimizer kernels can clear the gradient buffer after consuming it
        self._skip_zero = self.cuda and bool(getattr(inner, 'zero_grads_in_step', False))

    # -- one eager step ------------------------------------------------------
    def _loss(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        if self.cuda:
            from ..ops import fused
            loss = fused.try_loss(self.model, x, y, self.loss_name)
            if loss is not None:
                return loss
        return self.loss_fn(self.model(x), y)

    def _maybe_direct(self, x: torch.Tensor):
        if self._direct_checked or not self.use_direct:
            return self._direct
        self._direct_checked = True
        from ..ops.fused import DirectStep
        if DirectStep.supported(self.model, self.optimizer, self.loss_name, x):
            self._direct = DirectStep(self.model, self.optimizer, self.loss_name)
            self._skip_zero = True
        return self._direct

    def _eager_step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        direct = self._maybe_direct(x)
        if direct is not None:
            loss = direct.run(x, y)
            self.optimizer.step()
            return loss.reshape(())
        if not self._skip_zero:
            self.optimizer.zero_grad()
        loss = self._loss(x, y)
        if self.cuda:
            from ..ops.heads import unit_seed
            with unit_seed():
                loss.backward()
        else:
            loss.backward()
        self.optimizer.step()
        return loss.detach()

    # -- graph capture -------------------------------------------------------
    def _capture(self, x: torch.Tensor, y: torch.Tensor) -> None:
        self._static_x = x.clone()
        self._static_y = y.clone()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._static_loss = self._eager_step(self._static_x, self._static_y)
        self._graph = g

    def notify_params_changed(self) -> None:
        """Call after changing parameters outside the optimizer (e.g. an ADAG pull): refreshes bf16 shadows."""
        if hasattr(self._inner, 'refresh_shadows'):
            self._inner.refresh_shadows()

    def step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        self.steps += 1
        if self.cuda and hasattr(self._inner, 'refresh_shadows_if_stale'):
            self._inner.refresh_shadows_if_stale()
        if not self.use_cuda_graph:
            return self._eager_step(x, y)
        if self._graph is None:
            if self._warm < 3:  # warm up allocators / lazy inits on a side stream before capturing
                self._warm += 1
                s = torch.cuda.Stream(device=self.device)
                s.wait_stream(torch.cuda.current_stream(self.device))
                with torch.cuda.stream(s):
                    loss = self._eager_step(x, y)
                torch.cuda.current_stream(self.device).wait_stream(s)
                return loss
            self._capture(x, y)
            # the capture itself does not run the step: replay below with this batch
        if x.shape != self._static_x.shape:
            return self._eager_step(x, y)
        self._static_x.copy_(x, non_blocking=True)
        self._static_y.copy_(y, non_blocking=True)
        self._graph.replay()
        return self._static_loss

    @torch.no_grad()
    def predict(self, x: torch.Tensor) -> torch.Tensor:
        return self.model(x)
