"""Pattern dispatch: recognise the reference's three architectures inside a ``Sequential`` and run them
through the fused sm_100a kernels (recurrent layer + fused head; CNN trunk + fused head)."""
from __future__ import annotations

from typing import Optional

import torch

from ..models import layers as L


def _plan(model):
    plan = getattr(model, '_fused_plan', None)
    if plan is not None:
        return plan
    ls = list(model.layers_list)
    kinds = [l.keras_class for l in ls]
    plan = {'kind': None}
    if len(ls) == 2 and kinds[0] in ('LSTM', 'GRU') and kinds[1] == 'Dense' and ls[1].activation in (None, 'linear') \
            and ls[0].units == 128 and not ls[0].return_sequences and model.input_shape[-1] == 1 \
            and ls[1].units <= 4:
        plan = {'kind': 'rnn_regressor', 'rnn': ls[0], 'dense': ls[1]}
    elif kinds == ['Conv2D', 'Activation', 'Conv2D', 'Activation', 'MaxPooling2D', 'Flatten', 'Dense', 'Activation',
                   'Dense', 'Activation'] \
            and ls[1].activation == 'relu' and ls[3].activation == 'relu' and ls[7].activation == 'relu' \
            and ls[9].activation == 'softmax' and ls[8].units <= 32:
        plan = {'kind': 'cnn', 'conv1': ls[0], 'conv2': ls[2], 'dense1': ls[6], 'dense2': ls[8]}
    elif len(ls) >= 2 and kinds[-1] == 'Dense' and all(k == 'LSTM' for k in kinds[:-1]) \
            and ls[-1].activation in (None, 'linear') and ls[-1].units <= 4 and model.input_shape[-1] == 1 \
            and all(l.units % 128 == 0 for l in ls[:-1]) and all(l.return_sequences for l in ls[:-2]) \
            and not ls[-2].return_sequences and (len(ls) > 2 or ls[0].units > 128):
        plan = {'kind': 'stacked_lstm', 'lstms': ls[:-1], 'dense': ls[-1]}
    model._fused_plan = plan
    return plan


class NativePathError(RuntimeError):
    """A CUDA model would train on PyTorch's eager cuDNN / cuBLAS kernels because no hand-written kernel covers it."""


def covers(model, x: torch.Tensor, loss_name: str) -> bool:
    """Do the hand-written kernels cover training ``model`` on batch ``x`` with this loss (through the autograd-free
    DirectStep or the autograd wrappers of the same kernels)?"""
    plan = _plan(model)
    if not x.is_cuda:
        return False
    if plan['kind'] == 'rnn_regressor':
        from . import recurrent
        return recurrent.native_available(x, plan['rnn'].units)
    if plan['kind'] == 'cnn':
        from . import cnn
        return loss_name == 'categorical_crossentropy' and cnn.native_available(plan, x)
    if plan['kind'] == 'stacked_lstm':
        from . import lstm_big
        return loss_name in ('mean_squared_error', 'mse') and lstm_big.supported(plan['lstms'], x)
    return False


def require_native(model, x: torch.Tensor, loss_name: str) -> None:
    """Raise instead of silently training a CUDA model on eager PyTorch ops (set DDLS_ALLOW_EAGER=1 to allow it)."""
    import os
    if x.is_cuda and not covers(model, x, loss_name) and os.environ.get('DDLS_ALLOW_EAGER', '0') != '1':
        kinds = [l.keras_class for l in model.layers_list]
        raise NativePathError(
            f'no hand-written sm_100a kernel path covers this model on CUDA (layers {kinds}, input {tuple(x.shape)}, loss '
            f'{loss_name!r}).  Covered: LSTM/GRU(128) + Dense(<=4) on [B,T,1] inputs with mean_squared_error; the 28x28x1 '
            f'Conv(32)-Conv(32)-MaxPool-Dense-Dense CNN with categorical_crossentropy; stacked LSTM (units % 128 == 0) + '
            f'Dense.  Set DDLS_ALLOW_EAGER=1 to train it on PyTorch eager kernels (cuDNN / cuBLAS) instead.')


def _cnn_trunk(plan, x):
    from . import cnn
    if cnn.native_available(plan, x):
        return cnn.trunk(plan, x)
    return None


def try_forward(model, x: torch.Tensor) -> Optional[torch.Tensor]:
    plan = _plan(model)
    if plan['kind'] == 'rnn_regressor':
        from . import heads
        h = plan['rnn'](x)
        if h.is_cuda and not torch.is_grad_enabled():
            return heads.dense_linear(h, plan['dense'].kernel, plan['dense'].bias)
        return plan['dense'](h)
    if plan['kind'] == 'stacked_lstm' and not torch.is_grad_enabled():
        from .. import _native as N
        from . import heads, lstm_big
        if lstm_big.supported(plan['lstms'], x):
            B, T, _ = x.shape
            key = ('stacked_infer', B, T)
            cache = model.__dict__.setdefault('_fused_cache', {})
            bufs = cache.get(key)
            if bufs is None:
                bufs = cache[key] = lstm_big.StackedLSTMBuffers(plan['lstms'], B, T, x.device)
            st = N.stream_ptr(x.device)
            lstm_big.prepare_weights(plan['lstms'], bufs, st)
            h = lstm_big.forward(plan['lstms'], bufs, x.reshape(B, T).contiguous(), st, save=False)
            return heads.dense_linear(h, plan['dense'].kernel, plan['dense'].bias)
    if plan['kind'] == 'cnn' and not torch.is_grad_enabled():
        from . import heads
        a = _cnn_trunk(plan, x)
        if a is not None:
            return heads.dense_softmax(a, plan['dense2'].kernel, plan['dense2'].bias)
    return None


def try_loss(model, x: torch.Tensor, y: torch.Tensor, loss_name: str) -> Optional[torch.Tensor]:
    plan = _plan(model)
    from . import heads
    if plan['kind'] == 'rnn_regressor' and loss_name in ('mean_squared_error', 'mse'):
        h = plan['rnn'](x)
        loss, _ = heads.dense_mse(h, plan['dense'].kernel, plan['dense'].bias, y)
        return loss
    if plan['kind'] == 'cnn' and loss_name == 'categorical_crossentropy':
        a = _cnn_trunk(plan, x)
        if a is None:   # trunk through torch ops, head still fused
            ls = list(model.layers_list)
            a = x
            for l in ls[:8]:
                a = l(a)
        loss, _ = heads.dense_softmax_xent(a, plan['dense2'].kernel, plan['dense2'].bias, y)
        return loss
    return None


# ----------------------------------------------------------------------------------------------------------
# Autograd-free training step
# ----------------------------------------------------------------------------------------------------------
class DirectStep:
    """Forward + backward of a recognised architecture as a fixed kernel sequence, with every gradient written
    straight into its slice of the flat gradient buffer (which is the allreduce buffer): no autograd graph,
    no gradient-accumulation kernels, no per-step allocations.  Requires a flat optimizer that clears the
    gradient buffer after consuming it (a few small gradients are accumulated with atomics)."""

    def __init__(self, model, optimizer, loss_name: str):
        from ..optim.keras_optimizers import _FlatOptimizer
        self.model = model
        self.inner = getattr(optimizer, 'inner', optimizer)
        if not isinstance(self.inner, _FlatOptimizer):
            raise TypeError('DirectStep needs a flat fused optimizer')
        self.plan = _plan(model)
        self.kind = self.plan['kind']
        self.loss_name = loss_name
        self.flat = self.inner.flat
        self.inner.zero_grads_in_step = True
        self.flat.zero_grads()
        if not self.flat.grads_attached():
            self.flat.attach_grads()
        self._bufs = {}
        self.dev = self.flat.device
        # called with each parameter as soon as its gradient is final, so a DistributedOptimizer can launch the
        # bucket's allreduce on its side stream while the rest of the backward kernels are still running
        self.grad_ready = getattr(optimizer, '_on_grad', None) if hasattr(optimizer, 'buckets') else None
        if self.kind == 'cnn':
            d1 = self.plan['dense1']
            self.ld1 = (d1.kernel.shape[1] + 7) // 8 * 8
            self.w1s = self.inner.add_shadow(d1.kernel, self.ld1)

    @staticmethod
    def supported(model, optimizer, loss_name: str, x: torch.Tensor) -> bool:
        from ..optim.keras_optimizers import _FlatOptimizer
        inner = getattr(optimizer, 'inner', optimizer)
        if not (x.is_cuda and isinstance(inner, _FlatOptimizer)):
            return False
        plan = _plan(model)
        if plan['kind'] == 'rnn_regressor' and loss_name in ('mean_squared_error', 'mse'):
            from . import recurrent
            return recurrent.native_available(x, plan['rnn'].units)
        if plan['kind'] == 'cnn' and loss_name == 'categorical_crossentropy':
            from . import cnn
            return cnn.native_available(plan, x) and plan['dense1'].kernel.shape[1] <= 256 and x.shape[0] <= 1024
        if plan['kind'] == 'stacked_lstm' and loss_name in ('mean_squared_error', 'mse'):
            from . import lstm_big
            return lstm_big.supported(plan['lstms'], x)
        return False

    def _ready(self, *params):
        if self.grad_ready is not None:
            for p in params:
                self.grad_ready(p)

    def _buf(self, name, shape, dtype):
        key = (name, tuple(shape), dtype)
        t = self._bufs.get(key)
        if t is None:
            t = torch.empty(shape, dtype=dtype, device=self.dev)
            self._bufs[key] = t
        return t

    def run(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        if self.kind == 'cnn':
            return self._run_cnn(x, y)
        if self.kind == 'stacked_lstm':
            return self._run_stacked(x, y)
        return self._run_rnn(x, y)

    # ------------------------------------------------------------------ stacked large-hidden LSTM
    def _run_stacked(self, x, y):
        from .. import _native as N
        from . import heads, lstm_big
        lstms, dense = self.plan['lstms'], self.plan['dense']
        B, T, _ = x.shape
        st = N.stream_ptr(self.dev)
        key = ('stacked', B, T)
        bufs = self._bufs.get(key)
        if bufs is None:
            bufs = self._bufs[key] = lstm_big.StackedLSTMBuffers(lstms, B, T, self.dev)
        x2 = x.reshape(B, T).contiguous()
        O = dense.kernel.shape[1]
        Hl = lstms[-1].units
        pred = self._buf('pred', (B, O), torch.float32)
        loss = self._buf('loss', (1,), torch.float32)
        dh = self._buf('dh', (B, Hl), torch.float32)
        lstm_big.prepare_weights(lstms, bufs, st)
        h = lstm_big.forward(lstms, bufs, x2, st, save=True)
        heads.dense_mse_raw(h, dense.kernel, dense.bias, y.reshape(B, O).contiguous(), pred, loss, dh, dense.kernel.grad,
                            dense.bias.grad)
        self._ready(dense.bias, dense.kernel)
        lstm_big.backward(lstms, bufs, x2, dh, st)
        for l in reversed(lstms):
            self._ready(l.bias, l.recurrent_kernel, l.kernel)
        return loss

    # ------------------------------------------------------------------ CNN
    def _run_cnn(self, x, y):
        from .. import _native as N
        from . import cnn as C, heads
        from .gemm import gemm_bf16, workspace
        p = self.plan
        c1, c2, d1, d2 = p['conv1'], p['conv2'], p['dense1'], p['dense2']
        B = x.shape[0]
        F = d1.kernel.shape[1]
        ld = self.ld1
        st = N.stream_ptr(self.dev)
        bf, f32 = torch.bfloat16, torch.float32
        a1 = self._buf('a1', (B, 26, 26, 32), bf)
        pooled = self._buf('pooled', (B, 12, 12, 32), bf)
        pidx = self._buf('pidx', (B, 12, 12, 32), torch.uint8)
        a3 = self._buf('a3', (B, F), f32)
        loss = self._buf('loss', (1,), f32)
        dz3 = self._buf('dz3', (B, ld), bf)
        dflat = self._buf('dflat', (B, 4608), bf)
        dz2 = self._buf('dz2', (B, 24, 24, 32), bf)
        dz1 = self._buf('dz1', (B, 26, 26, 32), bf)
        x = x.contiguous()
        N.call('ddls_cnn_conv1_fwd', x.data_ptr(), c1.kernel.data_ptr(), c1.bias.data_ptr(), a1.data_ptr(), B, st)
        N.call(C.conv2_fwd_name(), a1.data_ptr(), c2.kernel.data_ptr(), c2.bias.data_ptr(), pooled.data_ptr(),
               pidx.data_ptr(), B, st)
        flat = pooled.view(B, 4608)
        w1s = self.w1s[:, :F]
        gemm_bf16(flat, w1s, bias=d1.bias, relu=True, b_mn_major=True, split_k=C.FWD_SPLIT_K, out=a3)
        scr = self._bufs.get(('head_scratch', B))
        if scr is None:
            scr = self._bufs[('head_scratch', B)] = heads.softmax_xent_scratch(B, self.dev)
        heads.softmax_xent_raw(a3, d2.kernel, d2.bias, y.contiguous(), None, loss, None, dz3, ld, d2.kernel.grad,
                               d2.bias.grad, d1.bias.grad, True, scr)
        self._ready(d2.bias, d2.kernel, d1.bias)
        # data gradient first: it is the last reader of the bf16 shadow of dense1's kernel, and the bucket kernel launched
        # by _ready() below may rewrite that shadow (sharded step: reduce-scatter + optimizer + all-gather) on any rank
        gemm_bf16(dz3[:, :F], w1s, out=dflat)
        gemm_bf16(flat, dz3[:, :F], a_mn_major=True, b_mn_major=True, out=d1.kernel.grad)
        self._ready(d1.kernel)                    # 99 % of the gradient bytes: its bucket overlaps the conv backward
        N.call('ddls_cnn_unpool_bwd', dflat.data_ptr(), pooled.data_ptr(), pidx.data_ptr(), dz2.data_ptr(),
               c2.bias.grad.data_ptr(), B, st)
        ws = workspace(self.dev, max(N.lib().ddls_cnn_conv2_wgrad_scratch_bytes(B), B * 320 * 4))
        N.call(C.conv2_wgrad_name(), a1.data_ptr(), dz2.data_ptr(), c2.kernel.grad.data_ptr(), ws.data_ptr(),
               ws.numel(), B, st)
        N.call(C.conv2_dgrad_name(), dz2.data_ptr(), c2.kernel.data_ptr(), a1.data_ptr(), dz1.data_ptr(), B, st)
        g1k, g1b = c1.kernel.grad, c1.bias.grad
        if g1b.data_ptr() == g1k.data_ptr() + 288 * 4:
            N.call('ddls_cnn_conv1_wgrad', x.data_ptr(), dz1.data_ptr(), g1k.data_ptr(), ws.data_ptr(), ws.numel(), B, st)
        else:
            g1 = self._buf('g1', (320,), f32)
            N.call('ddls_cnn_conv1_wgrad', x.data_ptr(), dz1.data_ptr(), g1.data_ptr(), ws.data_ptr(), ws.numel(), B, st)
            g1k.view(-1).copy_(g1[:288]); g1b.copy_(g1[288:])
        self._ready(c2.bias, c2.kernel, c1.bias, c1.kernel)
        return loss

    # ------------------------------------------------------------------ LSTM / GRU regressor
    def _run_rnn(self, x, y):
        from .. import _native as N
        from . import heads
        from . import recurrent as R
        p = self.plan
        rnn, dense = p['rnn'], p['dense']
        is_lstm = rnn.keras_class == 'LSTM'
        H = R.H
        B, T, _ = x.shape
        bf, f32 = torch.bfloat16, torch.float32
        x2 = x.reshape(B, T).contiguous()
        O = dense.kernel.shape[1]
        h = self._buf('h', (B, H), f32)
        pred = self._buf('pred', (B, O), f32)
        loss = self._buf('loss', (1,), f32)
        dh = self._buf('dh', (B, H), f32)
        k, r, b = rnn.kernel, rnn.recurrent_kernel, rnn.bias
        if is_lstm:
            key = ('lstm_v2', B, T)
            bufs = self._bufs.get(key)
            if bufs is None:
                bufs = self._bufs[key] = R.lstm_v2_buffers(B, T, self.dev) + (
                    torch.empty(18 * 64 * 64, dtype=bf, device=self.dev),)
            hs2, gates2, cs2, dz2, img = bufs
            R.pack_image(r, k, b, 4, out=img)
            R.lstm_v2_forward(x2, img, h, hs2, gates2, cs2, B, T, False, True)
            heads.dense_mse_raw(h, dense.kernel, dense.bias, y.reshape(B, O).contiguous(), pred, loss, dh,
                                dense.kernel.grad, dense.bias.grad)
            R.lstm_v2_backward(dh, x2, img, gates2, cs2, dz2, k.grad, b.grad, r.grad, B, T, False, hs2)
        else:
            key = ('gru_v2', B, T)
            bufs = self._bufs.get(key)
            if bufs is None:
                bufs = self._bufs[key] = R.gru_v2_buffers(B, T, self.dev) + (
                    torch.empty(18 * 48 * 64, dtype=bf, device=self.dev),)
            hs2, rh2, zrh2, hprev2, da2, img = bufs
            R.pack_image(r, k, b, 3, out=img)
            R.gru_v2_forward(x2, img, h, hs2, rh2, zrh2, hprev2, B, T, False, True)
            heads.dense_mse_raw(h, dense.kernel, dense.bias, y.reshape(B, O).contiguous(), pred, loss, dh,
                                dense.kernel.grad, dense.bias.grad)
            R.gru_v2_backward(dh, x2, img, zrh2, hprev2, da2, k.grad, b.grad, r.grad, B, T, False, hs2, rh2)
        self._ready(dense.bias, dense.kernel, b, r, k)
        return loss
