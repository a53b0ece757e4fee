"""Keras-2 form Adam / Adagrad / SGD over a flat parameter space, one fused launch per step.

The reference names its worker optimizers by string — ``'adam'``
(/root/reference/ddl_mnist_aztk.py:195, /root/reference/ddl_nyiso_aztk.py:255) and
``'adagrad'`` (ddl_nyiso_aztk.py:207) — so the update rules are Keras':

    Adam:    lr_t = lr*sqrt(1-b2^t)/(1-b1^t);  p -= lr_t*m/(sqrt(v)+eps),  eps = 1e-7
    Adagrad: a += g^2;                         p -= lr*g/(sqrt(a)+eps),    lr = 0.01

On CUDA the whole step is one hand-written kernel over the flat buffers
(``ops.optimizer``), with the step counter kept in device memory so a captured
CUDA graph replays correctly.  On CPU the same math runs as flat torch ops.
"""
from __future__ import annotations

import math
from typing import Dict, Iterable, List, Optional

import torch

from .flat import FlatSpace


class _FlatOptimizer:
    kind = 'base'

    def __init__(self, params: Iterable[torch.nn.Parameter], comm=None, flat: Optional[FlatSpace] = None,
                 metric_slots: int = 0):
        self.flat = flat if flat is not None else FlatSpace(list(params), comm=comm, metric_slots=metric_slots)
        self.device = self.flat.device
        self._step_buf = torch.zeros(2, dtype=torch.int32, device=self.device)  # [steps, kernel ticket]
        self.step_count = self._step_buf[:1]
        self.grad_scale = 1.0  # folded into the update kernel (e.g. 1/backward_passes)
        self.zero_grads_in_step = False  # fused kernels can clear gradients after consuming them
        self._native = None
        if self.device.type == 'cuda':
            from ..ops import optimizer as native
            self._native = native
        self._shadows = []          # [(begin, rows, cols, ld, bf16 tensor)]
        self._shadow_by_param = {}
        self._specs, self._nspecs = None, 0
        self._seen_param_version = self._param_version()
        self._shard_gather = None   # set by DistributedOptimizer when the sharded step kernel owns the slots

    def gather_slots(self) -> None:
        """Collective no-op unless a ``DistributedOptimizer`` shards the slots by rank (then: exchange the slices so
        every rank holds complete m / v / accumulator buffers, e.g. before a checkpoint)."""
        if self._shard_gather is not None:
            self._shard_gather()

    # bf16 shadow copies of weight matrices, refreshed inside the fused update kernel -------------------
    def add_shadow(self, param: torch.nn.Parameter, ld: Optional[int] = None) -> torch.Tensor:
        """Keep a bf16 copy of 2-D-viewed ``param`` ([rows, cols], rows padded to ``ld`` columns, ld % 8 == 0)
        up to date: the optimizer kernel rewrites it in the pass that updates the fp32 master."""
        key = id(param)
        if key in self._shadow_by_param:
            return self._shadow_by_param[key]
        if self._native is None:
            raise RuntimeError('shadow weights are a CUDA feature')
        idx = next(i for i, q in enumerate(self.flat.params) if q is param)
        cols = param.shape[-1]
        rows = param.numel() // cols
        ld = ld or (cols + 7) // 8 * 8
        t = torch.zeros((rows, ld), dtype=torch.bfloat16, device=self.device)
        self._shadows.append((self.flat.offsets[idx], rows, cols, ld, t))
        self._shadow_by_param[key] = t
        self._specs, self._nspecs = self._native.make_specs(self._shadows)
        self.refresh_shadows()
        return t

    def _param_version(self) -> int:
        # FlatSpace re-points every parameter with `p.data = view`, which gives each parameter its OWN version counter:
        # `p.copy_()` (Sequential.set_weights, load_checkpoint) does not bump the flat buffer's.  Track both.
        return self.flat.flat_params._version + sum(p._version for p in self.flat.params)

    def refresh_shadows(self) -> None:
        if self._nspecs:
            self._native.refresh_shadows(self.flat.flat_params, self._specs, self._nspecs)
        self._seen_param_version = self._param_version()

    def refresh_shadows_if_stale(self) -> None:
        """Parameters changed through torch ops since the last refresh (broadcast, set_weights, ADAG pull)?"""
        if self._nspecs and self._param_version() != self._seen_param_version:
            self.refresh_shadows()

    # torch.optim-like surface ----------------------------------------------
    @property
    def param_groups(self):
        return [{'params': self.flat.params, **self.hyper()}]

    def zero_grad(self, set_to_none: bool = False) -> None:
        self.flat.zero_grads()

    def hyper(self) -> Dict[str, float]:
        return {}

    def slots(self) -> Dict[str, torch.Tensor]:
        return {}

    def step(self) -> None:
        if not self.flat.grads_attached():
            self.flat.gather_stray_grads()
        self._update()

    def _update(self) -> None:  # pragma: no cover - interface
        raise NotImplementedError

    def can_update_range(self) -> bool:
        return self._native is not None and self.kind in ('adam', 'adagrad', 'sgd') and not getattr(self, 'momentum', 0.0)

    def update_range(self, lo: int, hi: int, advance: bool) -> None:
        """Apply the update to flat elements [lo, hi) only (CUDA): lets a caller update each gradient bucket as soon as
        its gradients are final instead of one pass over everything at the end of backward."""
        f = self.flat
        s1 = getattr(self, 'm', None) if self.kind == 'adam' else getattr(self, 'acc', None)
        self._native.step_range(self.kind, lo, hi, advance, f.flat_params, f.flat_grads, s1, getattr(self, 'v', None),
                                self.step_count, self.lr, getattr(self, 'beta_1', 0.0), getattr(self, 'beta_2', 0.0),
                                getattr(self, 'epsilon', 0.0), self.grad_scale, self.zero_grads_in_step, self._specs,
                                self._nspecs)

    # checkpoint -------------------------------------------------------------
    def state_dict(self) -> Dict:
        return {'kind': self.kind, 'hyper': self.hyper(), 'step': int(self.step_count.item()),
                'slots': {k: v.detach().cpu().clone() for k, v in self.slots().items()}}

    def load_state_dict(self, sd: Dict) -> None:
        if sd['kind'] != self.kind:
            raise ValueError(f"optimizer kind mismatch: {sd['kind']} vs {self.kind}")
        self.step_count.fill_(int(sd['step']))
        for k, v in self.slots().items():
            v.copy_(sd['slots'][k].to(v.device))


class KerasAdam(_FlatOptimizer):
    kind = 'adam'

    def __init__(self, params, lr=1e-3, beta_1=0.9, beta_2=0.999, epsilon=1e-7, comm=None, flat=None, metric_slots=0):
        super().__init__(params, comm=comm, flat=flat, metric_slots=metric_slots)
        self.lr, self.beta_1, self.beta_2, self.epsilon = lr, beta_1, beta_2, epsilon
        self.m = self.flat.new_slot()
        self.v = self.flat.new_slot()

    def hyper(self):
        return {'lr': self.lr, 'beta_1': self.beta_1, 'beta_2': self.beta_2, 'epsilon': self.epsilon}

    def slots(self):
        return {'m': self.m, 'v': self.v}

    def _update(self):
        f = self.flat
        if self._native is not None:
            self._native.adam_step(f.flat_params, f.flat_grads, self.m, self.v, self.step_count,
                                   self.lr, self.beta_1, self.beta_2, self.epsilon, self.grad_scale,
                                   self.zero_grads_in_step, self._specs, self._nspecs)
            return
        self.step_count += 1
        t = int(self.step_count.item())
        lr_t = self.lr * math.sqrt(1.0 - self.beta_2 ** t) / (1.0 - self.beta_1 ** t)
        fg = f.flat_grads[:f.numel]
        g = fg * self.grad_scale if self.grad_scale != 1.0 else fg
        self.m.mul_(self.beta_1).add_(g, alpha=1 - self.beta_1)
        self.v.mul_(self.beta_2).addcmul_(g, g, value=1 - self.beta_2)
        f.flat_params.addcdiv_(self.m, self.v.sqrt().add_(self.epsilon), value=-lr_t)


class KerasAdagrad(_FlatOptimizer):
    kind = 'adagrad'

    def __init__(self, params, lr=0.01, epsilon=1e-7, comm=None, flat=None, metric_slots=0):
        super().__init__(params, comm=comm, flat=flat, metric_slots=metric_slots)
        self.lr, self.epsilon = lr, epsilon
        self.acc = self.flat.new_slot()

    def hyper(self):
        return {'lr': self.lr, 'epsilon': self.epsilon}

    def slots(self):
        return {'accumulator': self.acc}

    def _update(self):
        f = self.flat
        if self._native is not None:
            self._native.adagrad_step(f.flat_params, f.flat_grads, self.acc, self.step_count,
                                      self.lr, self.epsilon, self.grad_scale, self.zero_grads_in_step,
                                      self._specs, self._nspecs)
            return
        self.step_count += 1
        fg = f.flat_grads[:f.numel]
        g = fg * self.grad_scale if self.grad_scale != 1.0 else fg
        self.acc.addcmul_(g, g)
        f.flat_params.addcdiv_(g, self.acc.sqrt().add_(self.epsilon), value=-self.lr)


class KerasSGD(_FlatOptimizer):
    """Keras SGD: ``v = momentum*v - lr*g; p += v`` (``nesterov``: ``p += momentum*v - lr*g``).  Without momentum the
    update is one fused kernel; with momentum (dist-keras' EAMSGD workers) it is three torch ops over the flat buffers."""
    kind = 'sgd'

    def __init__(self, params, lr=0.01, momentum=0.0, nesterov=False, comm=None, flat=None, metric_slots=0):
        super().__init__(params, comm=comm, flat=flat, metric_slots=metric_slots)
        self.lr, self.momentum, self.nesterov = lr, float(momentum), bool(nesterov)
        self.velocity = torch.zeros_like(self.flat.flat_params) if self.momentum else None

    def hyper(self):
        return {'lr': self.lr, 'momentum': self.momentum, 'nesterov': self.nesterov}

    def slots(self):
        return {'velocity': self.velocity} if self.velocity is not None else {}

    def _update(self):
        f = self.flat
        if self._native is not None and not self.momentum:
            self._native.sgd_step(f.flat_params, f.flat_grads, self.step_count, self.lr, self.grad_scale,
                                  self.zero_grads_in_step, self._specs, self._nspecs)
            return
        self.step_count += 1
        g = f.flat_grads[:f.numel]
        if not self.momentum:
            f.flat_params.add_(g, alpha=-self.lr * self.grad_scale)
        else:
            self.velocity.mul_(self.momentum).add_(g, alpha=-self.lr * self.grad_scale)
            if self.nesterov:
                f.flat_params.add_(self.velocity, alpha=self.momentum).add_(g, alpha=-self.lr * self.grad_scale)
            else:
                f.flat_params.add_(self.velocity)
        if self._native is not None:                  # CUDA without the fused kernel: honour the fused contracts
            if self.zero_grads_in_step:
                g.zero_()
            self.refresh_shadows()


_BY_NAME = {'adam': KerasAdam, 'adagrad': KerasAdagrad, 'sgd': KerasSGD}


def get_optimizer(name_or_opt, params, comm=None, **kw):
    """Resolve a Keras optimizer name (``'adam'``, ``'adagrad'``, ``'sgd'``) to a flat fused optimizer."""
    if isinstance(name_or_opt, _FlatOptimizer):
        return name_or_opt
    key = str(name_or_opt).lower()
    if key not in _BY_NAME:
        raise ValueError(f'unsupported optimizer {name_or_opt!r}; have {sorted(_BY_NAME)}')
    return _BY_NAME[key](params, comm=comm, **kw)
