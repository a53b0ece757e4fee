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
        elif self._wants_local_buckets(base):
            # a lone CUDA rank: the same bucket machinery updates each gradient bucket as soon as it is final, on a side
            # stream, instead of one optimizer pass after backward (DistributedOptimizer 'local' mode)
            from ..parallel import DistributedOptimizer
            wrapped = DistributedOptimizer(base)
            if getattr(wrapped, '_local', False):
                base = wrapped
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
        self._native_checked = False
        self.use_direct = self.cuda and env_flag('DDLS_DIRECT_STEP', True)
        # the fused optimizer kernels can clear the gradient buffer after consuming it
        self._skip_zero = self.cuda and bool(getattr(inner, 'zero_grads_in_step', False))

    def _wants_local_buckets(self, base) -> bool:
        import os
        from .. import runtime
        from ..optim.keras_optimizers import _FlatOptimizer
        if not (self.device.type == 'cuda' and isinstance(base, _FlatOptimizer)) or os.environ.get('DDLS_LOCAL_BUCKETS', '1') == '0':
            return False
        if runtime.is_initialized():
            return runtime.size() == 1
        return int(os.environ.get('WORLD_SIZE', '1')) == 1 and int(os.environ.get('DDLS_SIZE', '1')) == 1

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
        if self.cuda and not self._native_checked:
            from ..ops import fused
            fused.require_native(self.model, x, self.loss_name)      # no silent cuDNN / cuBLAS training on a GPU box
            self._native_checked = True
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
        if self._static_x is None or self._static_x.shape != x.shape:
            self._static_x = x.clone()
            self._static_y = y.clone()
        elif x.data_ptr() != self._static_x.data_ptr():
            self._static_x.copy_(x); self._static_y.copy_(y)
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
        if x.data_ptr() != self._static_x.data_ptr():      # callers may fill input_buffers() directly (no extra hop)
            self._static_x.copy_(x, non_blocking=True)
        if y.data_ptr() != self._static_y.data_ptr():
            self._static_y.copy_(y, non_blocking=True)
        self._graph.replay()
        return self._static_loss

    def input_buffers(self, x_like: torch.Tensor, y_like: torch.Tensor):
        """The step's static input tensors (created from the given shapes/dtypes on first use).  A feeder that copies
        its pinned host batch straight into them and then calls ``step(xb, yb)`` with these very tensors saves the
        device-to-device hop a captured graph otherwise needs."""
        if self._static_x is None:
            self._static_x = torch.empty(x_like.shape, dtype=x_like.dtype, device=self.device)
            self._static_y = torch.empty(y_like.shape, dtype=y_like.dtype, device=self.device)
        return self._static_x, self._static_y

    # -- pipelined host feed ------------------------------------------------------------------------------------
    def run_stream(self, batches, x_like: torch.Tensor, y_like: torch.Tensor, stage=None):
        """Train on an iterable of pinned HOST batches with the copies pipelined against the compute:

        * two input slots, each with its own captured step graph; the host->device copy (and any on-device
          preprocessing done by ``stage``) of batch i+1 runs on a copy stream while the graph of batch i executes,
        * the loss of every step is copied back to pinned host memory, but the host only waits for the loss of step
          i-1 after it has launched step i — the launch latencies of one step hide behind the execution of another.

        ``stage(host_batch, x_slot, y_slot)`` fills the slot's device tensors (default: ``x_slot.copy_(hb[0]);
        y_slot.copy_(hb[1])``); it runs with the copy stream current.  ``x_like``/``y_like`` give the shape and dtype
        of the model inputs.  Host batches must stay alive until their step has run.  Returns the per-step losses."""
        if not self.cuda:
            out = []
            for hb in batches:
                out.append(float(self._eager_step(hb[0], hb[1])))
            return out
        if getattr(self, '_slots', None) is None:
            self._slots = [dict(x=torch.empty(x_like.shape, dtype=x_like.dtype, device=self.device),
                                y=torch.empty(y_like.shape, dtype=y_like.dtype, device=self.device),
                                graph=None, loss=None, ready=torch.cuda.Event(), free=torch.cuda.Event(),
                                done=torch.cuda.Event(), loss_host=torch.empty(1, dtype=torch.float32).pin_memory())
                           for _ in range(2)]
            self._copy_stream = torch.cuda.Stream(device=self.device)
        if stage is None:
            def stage(hb, xs, ys):
                xs.copy_(hb[0], non_blocking=True); ys.copy_(hb[1], non_blocking=True)
        cur = torch.cuda.current_stream(self.device)
        losses, prev = [], None
        for i, hb in enumerate(batches):
            sl = self._slots[i & 1]
            with torch.cuda.stream(self._copy_stream):
                if i >= 2:
                    self._copy_stream.wait_event(sl['free'])     # the step two batches ago has consumed this slot
                stage(hb, sl['x'], sl['y'])
                sl['ready'].record(self._copy_stream)
            cur.wait_event(sl['ready'])
            loss = self._slot_step(sl)
            sl['free'].record(cur)
            sl['loss_host'].copy_(loss.reshape(1), non_blocking=True)
            sl['done'].record(cur)
            if prev is not None:
                prev['done'].synchronize()
                losses.append(float(prev['loss_host'][0]))
            prev = sl
        if prev is not None:
            prev['done'].synchronize()
            losses.append(float(prev['loss_host'][0]))
        return losses

    def _slot_step(self, sl) -> torch.Tensor:
        self.steps += 1
        if hasattr(self._inner, 'refresh_shadows_if_stale'):
            self._inner.refresh_shadows_if_stale()
        if not self.use_cuda_graph:
            return self._eager_step(sl['x'], sl['y'])
        if sl['graph'] is None:
            if self._warm < 3:
                self._warm += 1
                s = torch.cuda.Stream(device=self.device)
                s.wait_stream(torch.cuda.current_stream(self.device))
                with torch.cuda.stream(s):
                    loss = self._eager_step(sl['x'], sl['y'])
                torch.cuda.current_stream(self.device).wait_stream(s)
                return loss
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                sl['loss'] = self._eager_step(sl['x'], sl['y'])
            sl['graph'] = g
        sl['graph'].replay()
        return sl['loss']

    # -- device-resident dataset ------------------------------------------------------------------------------
    def attach_dataset(self, X: torch.Tensor, Y: torch.Tensor, batch_size: int, perm: Optional[torch.Tensor] = None) -> None:
        """Keep the (partition of the) training set on the device — the counterpart of the reference's cached
        DataFrame partition (/root/reference/ddl_mnist_aztk.py:156-161).  ``step_dataset()`` then trains on
        consecutive minibatches (positions wrap around the end like dist-keras' batching across epochs); the
        minibatch gather is a kernel inside the captured step, driven by a device-side cursor, so a step is ONE graph
        launch with no input copy.  ``perm`` (int32 [N]) gives a shuffled order."""
        from .. import _native as N
        assert self.cuda and X.is_cuda and Y.is_cuda and X.shape[0] == Y.shape[0]
        X, Y = X.contiguous(), Y.contiguous()
        rx, ry = X[0].numel() * X.element_size(), Y[0].numel() * Y.element_size()
        if rx % 4 or ry % 4:
            raise ValueError('dataset rows must be multiples of 4 bytes')
        N.register('ddls_gather_batch', [N.c_void_p, N.c_void_p, N.c_void_p, N.c_void_p, N.c_void_p, N.c_void_p,
                                         N.c_int, N.c_longlong, N.c_longlong, N.c_longlong, N.c_void_p])
        self._ds = dict(X=X, Y=Y, B=int(batch_size), rx=rx, ry=ry,
                        perm=None if perm is None else perm.to(self.device, torch.int32).contiguous(),
                        cursor=torch.zeros(2, dtype=torch.int32, device=self.device))
        self._ds_x = torch.empty((batch_size,) + tuple(X.shape[1:]), dtype=X.dtype, device=self.device)
        self._ds_y = torch.empty((batch_size,) + tuple(Y.shape[1:]), dtype=Y.dtype, device=self.device)
        self._ds_graph = None
        self._ds_warm = 0

    def _gather(self) -> None:
        from .. import _native as N
        d = self._ds
        N.call('ddls_gather_batch', d['X'].data_ptr(), d['Y'].data_ptr(), self._ds_x.data_ptr(), self._ds_y.data_ptr(),
               N.ptr(d['perm']), d['cursor'].data_ptr(), d['B'], d['X'].shape[0], d['rx'], d['ry'],
               N.stream_ptr(self.device))

    def step_dataset(self) -> torch.Tensor:
        """One training step on the next minibatch of the attached dataset."""
        self.steps += 1
        if hasattr(self._inner, 'refresh_shadows_if_stale'):
            self._inner.refresh_shadows_if_stale()
        if self.use_cuda_graph and self._ds_graph is not None:
            self._ds_graph.replay()
            return self._ds_loss
        if not self.use_cuda_graph or self._ds_warm < 3:
            self._ds_warm += 1
            s = torch.cuda.Stream(device=self.device)
            s.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(s):
                self._gather()
                loss = self._eager_step(self._ds_x, self._ds_y)
            torch.cuda.current_stream(self.device).wait_stream(s)
            return loss
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._gather()
            self._ds_loss = self._eager_step(self._ds_x, self._ds_y)
        self._ds_graph = g
        g.replay()
        return self._ds_loss

    def dataset_cursor(self) -> int:
        """Number of minibatches consumed from the attached dataset (device -> host read)."""
        return int(self._ds['cursor'][0].item())

    @torch.no_grad()
    def predict(self, x: torch.Tensor) -> torch.Tensor:
        return self.model(x)
