"""Fused terminal heads: Dense + MSE and Dense + softmax + categorical cross-entropy, forward and
backward in ONE launch each (``csrc/head.cu``)."""
from __future__ import annotations

import torch

from .. import _native as N

P, I = N.c_void_p, N.c_int
N.register('ddls_dense_mse', [P, P, P, P, P, P, P, P, P, I, I, I, P])
N.register('ddls_dense_softmax_xent', [P, P, P, P, P, P, P, P, P, I, I, I, P])


class _DenseMSE(torch.autograd.Function):
    """loss = mean_b mean_o (h·W + b - y)^2 ; gradients are produced by the same kernel."""

    @staticmethod
    def forward(ctx, h, w, b, y):
        B, Hd = h.shape
        O = w.shape[1]
        dev = h.device
        h = h.contiguous(); y2 = y.reshape(B, O).contiguous()
        acc = torch.zeros(1 + Hd * O + O, dtype=torch.float32, device=dev)
        loss, dw, db = acc[0:1], acc[1:1 + Hd * O], acc[1 + Hd * O:]
        pred = torch.empty((B, O), dtype=torch.float32, device=dev)
        dh = torch.empty_like(h)
        N.call('ddls_dense_mse', h.data_ptr(), w.data_ptr(), b.data_ptr(), y2.data_ptr(), pred.data_ptr(),
               loss.data_ptr(), dh.data_ptr(), dw.data_ptr(), db.data_ptr(), B, Hd, O, N.stream_ptr(dev))
        ctx.save_for_backward(dh, dw.view(Hd, O), db)
        ctx.mark_non_differentiable(pred)
        return loss.reshape(()), pred

    @staticmethod
    def backward(ctx, gloss, _gpred):
        dh, dw, db = ctx.saved_tensors
        if _UNIT_SEED[0]:          # the step engine seeds loss.backward() with exactly 1
            return dh, dw, db, None
        return dh * gloss, dw * gloss, db * gloss, None


_UNIT_SEED = [False]


class unit_seed:
    """Context in which the caller guarantees ``loss.backward()`` is seeded with 1 (the step engine
    does), so the gradients the fused head already produced need no rescaling kernels."""

    def __enter__(self):
        self._prev = _UNIT_SEED[0]
        _UNIT_SEED[0] = True

    def __exit__(self, *a):
        _UNIT_SEED[0] = self._prev


def dense_mse(h, w, b, y):
    """Returns (loss, pred)."""
    return _DenseMSE.apply(h, w, b, y)


def dense_linear(h, w, b):
    """Inference-only Dense(H->O) through the same kernel (no targets)."""
    B, Hd = h.shape
    O = w.shape[1]
    pred = torch.empty((B, O), dtype=torch.float32, device=h.device)
    N.call('ddls_dense_mse', h.contiguous().data_ptr(), w.data_ptr(), b.data_ptr(), 0, pred.data_ptr(), 0, 0, 0, 0,
           B, Hd, O, N.stream_ptr(h.device))
    return pred


class _DenseSoftmaxXent(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, w, b, onehot):
        B, F = a.shape
        C = w.shape[1]
        dev = a.device
        a = a.contiguous(); onehot = onehot.contiguous()
        acc = torch.zeros(1 + F * C + C, dtype=torch.float32, device=dev)
        loss, dw, db = acc[0:1], acc[1:1 + F * C], acc[1 + F * C:]
        probs = torch.empty((B, C), dtype=torch.float32, device=dev)
        da = torch.empty_like(a)
        N.call('ddls_dense_softmax_xent', a.data_ptr(), w.data_ptr(), b.data_ptr(), onehot.data_ptr(),
               probs.data_ptr(), loss.data_ptr(), da.data_ptr(), dw.data_ptr(), db.data_ptr(), B, F, C,
               N.stream_ptr(dev))
        ctx.save_for_backward(da, dw.view(F, C), db)
        ctx.mark_non_differentiable(probs)
        return loss.reshape(()), probs

    @staticmethod
    def backward(ctx, gloss, _gp):
        da, dw, db = ctx.saved_tensors
        if _UNIT_SEED[0]:
            return da, dw, db, None
        return da * gloss, dw * gloss, db * gloss, None


def dense_softmax_xent(a, w, b, onehot):
    """Returns (loss, probs)."""
    return _DenseSoftmaxXent.apply(a, w, b, onehot)


def dense_softmax(a, w, b):
    B, F = a.shape
    C = w.shape[1]
    probs = torch.empty((B, C), dtype=torch.float32, device=a.device)
    N.call('ddls_dense_softmax_xent', a.contiguous().data_ptr(), w.data_ptr(), b.data_ptr(), 0, probs.data_ptr(), 0, 0,
           0, 0, B, F, C, N.stream_ptr(a.device))
    return probs
