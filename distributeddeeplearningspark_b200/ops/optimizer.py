"""Fused flat optimizer steps (one launch per step) — wrappers over ``csrc/elementwise.cu``."""
from __future__ import annotations

import torch

from .. import _native as N


def _step_ptr(step_count: torch.Tensor) -> int:
    # step_count is a view of a 2-element int32 buffer: [steps, arrival ticket]
    base = step_count._base if step_count._base is not None else step_count
    if base.numel() < 2:
        raise ValueError('step counter must be backed by a 2-element int32 buffer')
    return base.data_ptr()


def adam_step(p, g, m, v, step_count, lr, b1, b2, eps, grad_scale=1.0, zero_grads=False):
    N.call('ddls_adam_step', p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), _step_ptr(step_count),
           p.numel(), lr, b1, b2, eps, grad_scale, int(zero_grads), N.stream_ptr(p.device))


def adagrad_step(p, g, acc, step_count, lr, eps, grad_scale=1.0, zero_grads=False):
    N.call('ddls_adagrad_step', p.data_ptr(), g.data_ptr(), acc.data_ptr(), _step_ptr(step_count), p.numel(),
           lr, eps, grad_scale, int(zero_grads), N.stream_ptr(p.device))


def sgd_step(p, g, step_count, lr, grad_scale=1.0, zero_grads=False):
    N.call('ddls_sgd_step', p.data_ptr(), g.data_ptr(), _step_ptr(step_count), p.numel(), lr, grad_scale,
           int(zero_grads), N.stream_ptr(p.device))
