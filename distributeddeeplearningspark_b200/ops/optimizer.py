"""Fused flat optimizer steps (one launch per step) — wrappers over ``csrc/elementwise.cu``."""
from __future__ import annotations

import ctypes
from typing import List, Optional, Tuple

import torch

from .. import _native as N

MAX_SHADOWS = 4


class ShadowSpec(ctypes.Structure):
    _fields_ = [('begin', ctypes.c_ulonglong), ('rows', ctypes.c_uint), ('cols', ctypes.c_uint), ('ld', ctypes.c_uint),
                ('kind', ctypes.c_uint), ('dst', ctypes.c_void_p)]


P, I, Z, F = N.c_void_p, N.c_int, N.c_size_t, N.c_float
SP = ctypes.POINTER(ShadowSpec)
N.register('ddls_adam_step', [P, P, P, P, P, Z, F, F, F, F, F, I, SP, I, P])
N.register('ddls_adagrad_step', [P, P, P, P, Z, F, F, F, I, SP, I, P])
N.register('ddls_sgd_step', [P, P, P, Z, F, F, I, SP, I, P])
N.register('ddls_refresh_shadows', [P, Z, SP, I, P])
N.register('ddls_adam_step_range', [P, P, P, P, P, Z, Z, I, F, F, F, F, F, I, SP, I, P])
N.register('ddls_adagrad_step_range', [P, P, P, P, Z, Z, I, F, F, F, I, SP, I, P])
N.register('ddls_sgd_step_range', [P, P, P, Z, Z, I, F, F, I, SP, I, P])


def make_specs(shadows: List[Tuple[int, int, int, int, torch.Tensor]]):
    """shadows: [(begin, rows, cols, ld, bf16 tensor)] -> (ctypes array or None, count)."""
    if not shadows:
        return None, 0
    if len(shadows) > MAX_SHADOWS:
        raise ValueError(f'at most {MAX_SHADOWS} shadow matrices per optimizer')
    arr = (ShadowSpec * len(shadows))()
    for k, (begin, rows, cols, ld, t) in enumerate(shadows):
        arr[k].begin, arr[k].rows, arr[k].cols, arr[k].ld, arr[k].kind, arr[k].dst = begin, rows, cols, ld, 0, t.data_ptr()
    return arr, len(shadows)


def _step_ptr(step_count: torch.Tensor) -> int:
    # step_count is a view of a 2-element int32 buffer: [steps, arrival ticket]
    base = step_count._base if step_count._base is not None else step_count
    if base.numel() < 2:
        raise ValueError('step counter must be backed by a 2-element int32 buffer')
    return base.data_ptr()


def adam_step(p, g, m, v, step_count, lr, b1, b2, eps, grad_scale=1.0, zero_grads=False, specs=None, nspecs=0):
    N.call('ddls_adam_step', p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), _step_ptr(step_count),
           p.numel(), lr, b1, b2, eps, grad_scale, int(zero_grads), specs, nspecs, N.stream_ptr(p.device))


def adagrad_step(p, g, acc, step_count, lr, eps, grad_scale=1.0, zero_grads=False, specs=None, nspecs=0):
    N.call('ddls_adagrad_step', p.data_ptr(), g.data_ptr(), acc.data_ptr(), _step_ptr(step_count), p.numel(),
           lr, eps, grad_scale, int(zero_grads), specs, nspecs, N.stream_ptr(p.device))


def sgd_step(p, g, step_count, lr, grad_scale=1.0, zero_grads=False, specs=None, nspecs=0):
    N.call('ddls_sgd_step', p.data_ptr(), g.data_ptr(), _step_ptr(step_count), p.numel(), lr, grad_scale,
           int(zero_grads), specs, nspecs, N.stream_ptr(p.device))


def refresh_shadows(p, specs, nspecs):
    if nspecs:
        N.call('ddls_refresh_shadows', p.data_ptr(), p.numel(), specs, nspecs, N.stream_ptr(p.device))


def step_range(kind: str, lo: int, hi: int, advance: bool, p, g, s1, s2, step_count, lr, b1, b2, eps, grad_scale=1.0,
               zero_grads=False, specs=None, nspecs=0):
    """Keras update of flat elements [lo, hi) only (one gradient bucket); ``advance``: publish step+1 (the last range call
    of a training step).  ``p`` / ``g`` / ``s1`` / ``s2`` are the FULL flat buffers."""
    st, sp = N.stream_ptr(p.device), _step_ptr(step_count)
    ptr = lambda t: t[lo:].data_ptr()
    n = hi - lo
    if kind == 'adam':
        N.call('ddls_adam_step_range', ptr(p), ptr(g), ptr(s1), ptr(s2), sp, n, lo, int(advance), lr, b1, b2, eps, grad_scale,
               int(zero_grads), specs, nspecs, st)
    elif kind == 'adagrad':
        N.call('ddls_adagrad_step_range', ptr(p), ptr(g), ptr(s1), sp, n, lo, int(advance), lr, eps, grad_scale,
               int(zero_grads), specs, nspecs, st)
    else:
        N.call('ddls_sgd_step_range', ptr(p), ptr(g), sp, n, lo, int(advance), lr, grad_scale, int(zero_grads), specs,
               nspecs, st)
