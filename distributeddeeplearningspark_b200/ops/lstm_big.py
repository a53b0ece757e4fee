"""Large-hidden stacked LSTM (hidden % 128 == 0, e.g. 1024): per-time-step fused tcgen05 GEMM launches
(``csrc/lstm_big.cu``) + large GEMMs for the hoisted input projection and the weight gradients."""
from __future__ import annotations

from typing import List

import torch

from .. import _native as N
from .gemm import gemm_bf16

P, I = N.c_void_p, N.c_int
N.register('ddls_lstm_big_perm_cast', [P, P, I, I, P])
N.register('ddls_lstm_big_perm_vec', [P, P, I, P])
N.register('ddls_lstm_big_unperm', [P, P, I, I, I, P])
N.register('ddls_lstm_big_colsum', [P, P, P, P, I, I, P])
N.register('ddls_lstm_big_zx1', [P, P, P, P, I, I, I, P])
N.register('ddls_lstm_big_step_fwd2', [P, P, P, P, P, P, I, I, I, I, P])
N.register('ddls_lstm_big_seq_tiles_per_launch', [I, I])
N.register('ddls_lstm_big_seq_fwd', [P, P, P, P, P, P, P, I, I, I, P])
N.register('ddls_lstm_big_seq_bwd', [P, P, P, P, P, P, P, I, I, I, P])
N.register('ddls_lstm_big_step_bwd2', [P, P, P, P, P, P, P, I, I, I, I, P])


class StackedLSTMBuffers:
    """Per-(B, T) persistent activations of an L-layer stack: everything the backward pass re-reads."""

    def __init__(self, layers, B: int, T: int, dev):
        bf, f32 = torch.bfloat16, torch.float32
        self.B, self.T = B, T
        self.H = [l.units for l in layers]
        self.hs = [torch.zeros((T + 1, B, h), dtype=bf, device=dev) for h in self.H]      # hs[0] = 0
        self.cs = [torch.zeros((T + 1, B, h), dtype=f32, device=dev) for h in self.H]     # cs[0] = 0
        self.gates = [torch.empty((T, B, 4 * h), dtype=bf, device=dev) for h in self.H]
        self.dz = [torch.empty((T, B, 4 * h), dtype=bf, device=dev) for h in self.H]
        self.dc = [torch.empty((B, h), dtype=f32, device=dev) for h in self.H]
        self.zx = [torch.empty((T, B, 4 * h), dtype=bf, device=dev) for h in self.H]
        self.dx = [None] + [torch.empty((T, B, self.H[i - 1]), dtype=bf, device=dev) for i in range(1, len(layers))]
        self.u_perm = [torch.empty((h, 4 * h), dtype=bf, device=dev) for h in self.H]
        self.w_perm = [None] + [torch.empty((self.H[i - 1], 4 * self.H[i]), dtype=bf, device=dev)
                                for i in range(1, len(layers))]
        self.b_perm = [torch.empty(4 * h, dtype=f32, device=dev) for h in self.H]
        self.wih_perm = torch.empty(4 * self.H[0], dtype=f32, device=dev)
        self.h_last = torch.empty((B, self.H[-1]), dtype=f32, device=dev)
        self.gtmp = torch.empty((max(self.H), 4 * max(self.H)), dtype=f32, device=dev)   # permuted weight-gradient staging
        self.xw = torch.empty(T * B, dtype=f32, device=dev)
        self.sync = torch.zeros(32 * max(4, (B + 127) // 128), dtype=torch.int32, device=dev)   # step-barrier counters


def use_seq(h: int, backward: bool) -> bool:
    """Whole-sequence persistent kernels unless DDLS_LSTM_BIG_SEQ=0 or the layer is too wide for one wave of CTAs."""
    import os
    if os.environ.get('DDLS_LSTM_BIG_SEQ', '1') == '0':
        return False
    return int(N.lib().ddls_lstm_big_seq_tiles_per_launch(h, int(backward))) >= 1


def supported(layers, x: torch.Tensor) -> bool:
    ok = (x.is_cuda and x.dtype == torch.float32 and x.dim() == 3 and x.shape[-1] == 1 and len(layers) >= 1
          and all(l.keras_class == 'LSTM' and l.units % 128 == 0 for l in layers)
          and all(l.return_sequences for l in layers[:-1]) and not layers[-1].return_sequences)
    if ok:
        N.lib()
    return ok


def prepare_weights(layers, bufs: StackedLSTMBuffers, st: int) -> None:
    """fp32 Keras weights -> permuted bf16 operands (once per optimizer step)."""
    for i, l in enumerate(layers):
        h = l.units
        N.call('ddls_lstm_big_perm_cast', l.recurrent_kernel.data_ptr(), bufs.u_perm[i].data_ptr(), h, h, st)
        N.call('ddls_lstm_big_perm_vec', l.bias.data_ptr(), bufs.b_perm[i].data_ptr(), h, st)
        if i == 0:
            N.call('ddls_lstm_big_perm_vec', l.kernel.data_ptr(), bufs.wih_perm.data_ptr(), h, st)
        else:
            N.call('ddls_lstm_big_perm_cast', l.kernel.data_ptr(), bufs.w_perm[i].data_ptr(), l.kernel.shape[0], h, st)


def forward(layers, bufs: StackedLSTMBuffers, x2: torch.Tensor, st: int, save: bool = True) -> torch.Tensor:
    """x2: [B, T] fp32.  Returns the last layer's final hidden state [B, H] (fp32)."""
    B, T = bufs.B, bufs.T
    for i, l in enumerate(layers):
        h = l.units
        last_layer = i == len(layers) - 1
        if i == 0:
            N.call('ddls_lstm_big_zx1', x2.data_ptr(), bufs.wih_perm.data_ptr(), bufs.b_perm[0].data_ptr(),
                   bufs.zx[0].data_ptr(), B, T, h, st)
        else:
            hin = bufs.H[i - 1]
            gemm_bf16(bufs.hs[i - 1][1:].reshape(T * B, hin), bufs.w_perm[i], bias=bufs.b_perm[i], b_mn_major=True,
                      out=bufs.zx[i].view(T * B, 4 * h))
        hs, cs = bufs.hs[i], bufs.cs[i]
        if use_seq(h, False):
            N.call('ddls_lstm_big_seq_fwd', hs.data_ptr(), bufs.u_perm[i].data_ptr(), bufs.zx[i].data_ptr(), cs.data_ptr(),
                   bufs.gates[i].data_ptr() if save else 0, bufs.h_last.data_ptr() if last_layer else 0,
                   bufs.sync.data_ptr(), B, T, h, st)
            continue
        for t in range(T):
            N.call('ddls_lstm_big_step_fwd2', hs.data_ptr(), bufs.u_perm[i].data_ptr(), bufs.zx[i].data_ptr(),
                   cs.data_ptr(), bufs.gates[i].data_ptr() if save else 0,
                   bufs.h_last.data_ptr() if (last_layer and t == T - 1) else 0, B, T, h, t, st)
    return bufs.h_last


def backward(layers, bufs: StackedLSTMBuffers, x2: torch.Tensor, dh_last: torch.Tensor, st: int) -> None:
    """BPTT through the stack.  Parameter gradients are written/accumulated into ``param.grad`` (bias and the
    input-dim-1 kernel are accumulated with atomics, so their grads must be zero on entry)."""
    B, T = bufs.B, bufs.T
    L = len(layers)
    for i in reversed(range(L)):
        l = layers[i]
        h = l.units
        top = i == L - 1
        dz, gates, cs = bufs.dz[i], bufs.gates[i], bufs.cs[i]
        for t in (reversed(range(T)) if not use_seq(h, True) else ()):
            N.call('ddls_lstm_big_step_bwd2', dz.data_ptr(), bufs.u_perm[i].data_ptr(),
                   bufs.dx[i + 1][t].data_ptr() if not top else 0, dh_last.data_ptr() if (top and t == T - 1) else 0,
                   gates.data_ptr(), cs.data_ptr(), bufs.dc[i].data_ptr(), B, T, h, t, st)
        if use_seq(h, True):
            N.call('ddls_lstm_big_seq_bwd', dz.data_ptr(), bufs.u_perm[i].data_ptr(),
                   bufs.dx[i + 1].data_ptr() if not top else 0, dh_last.data_ptr() if top else 0, gates.data_ptr(),
                   cs.data_ptr(), bufs.sync.data_ptr(), B, T, h, st)
        dz2 = dz.view(T * B, 4 * h)
        g = bufs.gtmp.view(-1)[:h * 4 * h].view(h, 4 * h)
        # recurrent kernel: dU = sum_t h_{t-1}^T dZ_t
        gemm_bf16(bufs.hs[i][:T].reshape(T * B, h), dz2, a_mn_major=True, b_mn_major=True, out=g)
        N.call('ddls_lstm_big_unperm', g.data_ptr(), l.recurrent_kernel.grad.data_ptr(), h, h, 0, st)
        if i == 0:
            bufs.xw.view(T, B).copy_(x2.t())
            N.call('ddls_lstm_big_colsum', dz2.data_ptr(), bufs.xw.data_ptr(), l.bias.grad.data_ptr(),
                   l.kernel.grad.data_ptr(), T * B, h, st)
        else:
            N.call('ddls_lstm_big_colsum', dz2.data_ptr(), 0, l.bias.grad.data_ptr(), 0, T * B, h, st)
            hin = bufs.H[i - 1]
            xin = bufs.hs[i - 1][1:].reshape(T * B, hin)
            gw = bufs.gtmp.view(-1)[:hin * 4 * h].view(hin, 4 * h)
            gemm_bf16(xin, dz2, a_mn_major=True, b_mn_major=True, out=gw)
            N.call('ddls_lstm_big_unperm', gw.data_ptr(), l.kernel.grad.data_ptr(), hin, h, 0, st)
            # gradient towards the layer below: dX = dZ · W^T
            gemm_bf16(dz2, bufs.w_perm[i], out=bufs.dx[i].view(T * B, hin))
