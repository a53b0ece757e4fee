"""Fused terminal heads: Dense + MSE and Dense + softmax + categorical cross-entropy, forward and
backward in ONE launch each (``csrc/head.cu``).  Every output is overwritten by the kernel."""
from __future__ import annotations

import torch

from .. import _native as N

P, I = N.c_void_p, N.c_int
N.register('ddls_dense_mse', [P, P, P, P, P, P, P, P, P, I, I, I, P])
N.register('ddls_dense_softmax_xent', [P, P, P, P, P, P, P, P, I, P, P, P, P, I, I, I, I, P])

_UNIT_SEED = [False]


class unit_seed:
    """Context in which the caller guarantees ``loss.backward()`` is seeded with 1 (the step engine
    does), so the gradients the fused head already produced need no rescaling kernels."""

    def __enter__(self):
        self._prev = _UNIT_SEED[0]
        _UNIT_SEED[0] = True

    def __exit__(self, *a):
        _UNIT_SEED[0] = self._prev


def dense_mse_raw(h, w, b, y, pred, loss, dh, dw, db):
    """Direct launch: all tensors preallocated (``y``/gradients may be None for inference)."""
    B, Hd = h.shape
    N.call('ddls_dense_mse', h.data_ptr(), w.data_ptr(), b.data_ptr(), N.ptr(y), pred.data_ptr(), N.ptr(loss),
           N.ptr(dh), N.ptr(dw), N.ptr(db), B, Hd, w.shape[1], N.stream_ptr(h.device))


def softmax_xent_scratch(B: int, device) -> torch.Tensor:
    """Zero-initialised scratch for the head kernel (arrival ticket + per-CTA loss partials); reusable across calls."""
    return torch.zeros((B + 7) // 8 + 2, dtype=torch.float32, device=device)


def softmax_xent_raw(a, w, b, onehot, probs, loss, da_f32, da_bf16, ldz, dw, db, dprev_bias, relu_input, scratch=None):
    """dw / db / dprev_bias are accumulated into (must be zero on entry); everything else is overwritten."""
    B, F = a.shape
    N.call('ddls_dense_softmax_xent', a.data_ptr(), w.data_ptr(), b.data_ptr(), N.ptr(onehot), N.ptr(probs),
           N.ptr(loss), N.ptr(da_f32), N.ptr(da_bf16), int(ldz), N.ptr(dw), N.ptr(db), N.ptr(dprev_bias),
           N.ptr(scratch), B, F, w.shape[1], int(relu_input), N.stream_ptr(a.device))


class _DenseMSE(torch.autograd.Function):
    """loss = mean_b mean_o (h·W + b - y)^2 ; gradients are produced by the same kernel."""

    @staticmethod
    def forward(ctx, h, w, b, y):
        B, Hd = h.shape
        O = w.shape[1]
        dev = h.device
        h = h.contiguous(); y2 = y.reshape(B, O).contiguous()
        buf = torch.empty(1 + Hd * O + O, dtype=torch.float32, device=dev)
        loss, dw, db = buf[0:1], buf[1:1 + Hd * O].view(Hd, O), buf[1 + Hd * O:]
        pred = torch.empty((B, O), dtype=torch.float32, device=dev)
        dh = torch.empty_like(h)
        dense_mse_raw(h, w, b, y2, pred, loss, dh, dw, db)
        ctx.save_for_backward(dh, dw, db)
        ctx.mark_non_differentiable(pred)
        return loss.reshape(()), pred

    @staticmethod
    def backward(ctx, gloss, _gpred):
        dh, dw, db = ctx.saved_tensors
        if _UNIT_SEED[0]:          # the step engine seeds loss.backward() with exactly 1
            return dh, dw, db, None
        return dh * gloss, dw * gloss, db * gloss, None


def dense_mse(h, w, b, y):
    """Returns (loss, pred)."""
    return _DenseMSE.apply(h, w, b, y)


def dense_linear(h, w, b):
    """Inference-only Dense(H->O) through the same kernel (no targets)."""
    pred = torch.empty((h.shape[0], w.shape[1]), dtype=torch.float32, device=h.device)
    dense_mse_raw(h.contiguous(), w, b, None, pred, None, None, None, None)
    return pred


class _DenseSoftmaxXent(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, w, b, onehot):
        B, F = a.shape
        C = w.shape[1]
        dev = a.device
        a = a.contiguous(); onehot = onehot.contiguous()
        nscr = (B + 7) // 8 + 2
        buf = torch.zeros(1 + F * C + C + nscr, dtype=torch.float32, device=dev)
        loss, dw, db = buf[0:1], buf[1:1 + F * C].view(F, C), buf[1 + F * C:1 + F * C + C]
        probs = torch.empty((B, C), dtype=torch.float32, device=dev)
        da = torch.empty_like(a)
        softmax_xent_raw(a, w, b, onehot, probs, loss, da, None, 0, dw, db, None, False, buf[1 + F * C + C:])
        ctx.save_for_backward(da, dw, db)
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
    probs = torch.empty((a.shape[0], w.shape[1]), dtype=torch.float32, device=a.device)
    softmax_xent_raw(a.contiguous(), w, b, None, probs, None, None, None, 0, None, None, None, False)
    return probs
