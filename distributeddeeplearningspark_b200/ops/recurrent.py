"""Autograd wrappers for the persistent tcgen05 recurrent kernels (``csrc/rnn128.cu``)."""
from __future__ import annotations

import torch

from .. import _native as N

P = N.c_void_p
I = N.c_int
N.register('ddls_rnn128_pack', [P, P, P, P, I, P])
N.register('ddls_lstm128_forward_v2', [P, P, P, P, P, P, I, I, I, I, I, P])
N.register('ddls_lstm128_backward_v2', [P, P, P, P, P, P, P, P, I, I, I, I, P])
N.register('ddls_rnn128_wgrad', [P, P, P, I, I, I, I, I, P])

N.register('ddls_gru128_forward_v2', [P, P, P, P, P, P, P, I, I, I, I, I, P])
N.register('ddls_gru128_backward_v2', [P, P, P, P, P, P, P, P, I, I, I, I, P])

H = 128


def _tile() -> int:
    """Batch rows per CTA for the v2 kernels (0 = pick by batch size; DDLS_RNN_TILE forces 16/32/64)."""
    import os
    return int(os.environ.get('DDLS_RNN_TILE', '0'))


def pack_image(recurrent, kernel, bias, gates: int, out=None):
    """fp32 Keras weights -> bf16 UMMA image [18][gates*16][8][8] (see csrc/rnn128.cu)."""
    if out is None:
        out = torch.empty(18 * gates * 16 * 64, dtype=torch.bfloat16, device=recurrent.device)
    N.call('ddls_rnn128_pack', recurrent.data_ptr(), kernel.data_ptr(), bias.data_ptr(), out.data_ptr(), gates,
           N.stream_ptr(recurrent.device))
    return out


def kgroups_padded(T: int, B8: int) -> int:
    return ((T - 1) * B8 + 7) // 8 * 8


def lstm_v2_buffers(B: int, T: int, dev, zero: bool = True):
    """Stash buffers of the v2 kernels.  hs / dz carry 8 zeroed padding k-groups for the wgrad GEMM."""
    B8 = (B + 7) // 8
    mk = torch.zeros if zero else torch.empty
    hs = mk(((T * B8 + 8) * H * 8,), dtype=torch.bfloat16, device=dev)
    gates = torch.empty((T * B8 * 4 * H * 8,), dtype=torch.bfloat16, device=dev)
    cs = torch.empty((T * B8 * H * 8,), dtype=torch.float32, device=dev)
    dz = mk(((T * B8 + 8) * 4 * H * 8,), dtype=torch.bfloat16, device=dev)
    return hs, gates, cs, dz


def lstm_v2_forward(x2, img, out, hs, gates, cs, B, T, return_sequences, save):
    N.call('ddls_lstm128_forward_v2', x2.data_ptr(), img.data_ptr(), out.data_ptr(), N.ptr(hs), N.ptr(gates), N.ptr(cs),
           B, T, int(return_sequences), int(save), _tile(), N.stream_ptr(x2.device))


def lstm_v2_backward(dout, x2, img, gates, cs, dz, d_wih, d_bias, d_rec, B, T, return_sequences, hs):
    """BPTT + recurrent weight gradient.  d_wih / d_bias are accumulated into, d_rec [128,512] is overwritten."""
    N.call('ddls_lstm128_backward_v2', dout.data_ptr(), x2.data_ptr(), img.data_ptr(), gates.data_ptr(), cs.data_ptr(),
           dz.data_ptr(), d_wih.data_ptr(), d_bias.data_ptr(), B, T, int(return_sequences), _tile(),
           N.stream_ptr(x2.device))
    B8 = (B + 7) // 8
    if T > 1:
        b_img = dz.data_ptr() + B8 * 4 * H * 16          # k-groups of t >= 1
        N.call('ddls_rnn128_wgrad', hs.data_ptr(), b_img, d_rec.data_ptr(), kgroups_padded(T, B8), 4 * H, 4 * H, 0,
               d_rec.stride(0), N.stream_ptr(x2.device))
    else:
        d_rec.zero_()


def native_available(x: torch.Tensor, units: int) -> bool:
    """The persistent kernels cover the reference's recurrent layers: 128 units, input dim 1, fp32 CUDA input.
    On a CUDA tensor of that shape a missing native library is an error, not a fallback."""
    if not (x.is_cuda and units == H and x.dim() == 3 and x.shape[-1] == 1 and x.dtype == torch.float32):
        return False
    if x.shape[1] * 32 * 4 > 60 * 1024:   # x tile must fit next to the resident weights
        return False
    N.lib()  # raises NativeError if the .so is missing
    return True


def gru_v2_buffers(B: int, T: int, dev):
    B8 = (B + 7) // 8
    hs = torch.zeros(((T * B8 + 8) * H * 8,), dtype=torch.bfloat16, device=dev)
    rh = torch.zeros(((T * B8 + 8) * H * 8,), dtype=torch.bfloat16, device=dev)
    zrh = torch.empty((T * B8 * 3 * H * 8,), dtype=torch.bfloat16, device=dev)
    hprev = torch.empty((T * B8 * H * 8,), dtype=torch.float32, device=dev)
    da = torch.zeros(((T * B8 + 8) * 3 * H * 8,), dtype=torch.bfloat16, device=dev)
    return hs, rh, zrh, hprev, da


def gru_v2_forward(x2, img, out, hs, rh, zrh, hprev, B, T, return_sequences, save):
    N.call('ddls_gru128_forward_v2', x2.data_ptr(), img.data_ptr(), out.data_ptr(), N.ptr(hs), N.ptr(rh), N.ptr(zrh),
           N.ptr(hprev), B, T, int(return_sequences), int(save), _tile(), N.stream_ptr(x2.device))


def gru_v2_backward(dout, x2, img, zrh, hprev, da, d_wih, d_bias, d_rec, B, T, return_sequences, hs, rh):
    N.call('ddls_gru128_backward_v2', dout.data_ptr(), x2.data_ptr(), img.data_ptr(), zrh.data_ptr(), hprev.data_ptr(),
           da.data_ptr(), d_wih.data_ptr(), d_bias.data_ptr(), B, T, int(return_sequences), _tile(),
           N.stream_ptr(x2.device))
    B8 = (B + 7) // 8
    if T > 1:
        st = N.stream_ptr(x2.device)
        kg = kgroups_padded(T, B8)
        b_img = da.data_ptr() + B8 * 3 * H * 16
        # dU_z, dU_r: h_{t-1}^T · [d(az) | d(ar)] ;  dU_h: (r ⊙ h_{t-1})^T · d(ah)
        N.call('ddls_rnn128_wgrad', hs.data_ptr(), b_img, d_rec.data_ptr(), kg, 2 * H, 3 * H, 0, d_rec.stride(0), st)
        N.call('ddls_rnn128_wgrad', rh.data_ptr() + B8 * H * 16, b_img, d_rec.data_ptr() + 2 * H * 4, kg, H, 3 * H, 2 * H,
               d_rec.stride(0), st)
    else:
        d_rec.zero_()


class _GRU128v2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, kernel, recurrent, bias, return_sequences):
        B, T, _ = x.shape
        dev = x.device
        x2 = x.reshape(B, T).contiguous()
        need_bwd = any(ctx.needs_input_grad)
        out = torch.empty((B, T, H) if return_sequences else (B, H), dtype=torch.float32, device=dev)
        img = pack_image(recurrent, kernel, bias, 3)
        if need_bwd:
            hs, rh, zrh, hprev, da = gru_v2_buffers(B, T, dev)
        else:
            hs = rh = zrh = hprev = da = None
        gru_v2_forward(x2, img, out, hs, rh, zrh, hprev, B, T, return_sequences, need_bwd)
        if need_bwd:
            ctx.save_for_backward(x2, kernel, img, hs, rh, zrh, hprev, da)
            ctx.return_sequences = return_sequences
            ctx.dims = (B, T)
        return out

    @staticmethod
    def backward(ctx, dout):
        x2, kernel, img, hs, rh, zrh, hprev, da = ctx.saved_tensors
        B, T = ctx.dims
        dev = x2.device
        small = torch.zeros(6 * H, dtype=torch.float32, device=dev)
        d_wih, d_bias = small[:3 * H], small[3 * H:]
        d_rec = torch.empty((H, 3 * H), dtype=torch.float32, device=dev)
        gru_v2_backward(dout.contiguous(), x2, img, zrh, hprev, da, d_wih, d_bias, d_rec, B, T, ctx.return_sequences, hs, rh)
        dx = None
        if ctx.needs_input_grad[0]:
            B8 = (B + 7) // 8
            daf = da[:T * B8 * 3 * H * 8].view(T, B8, 3 * H, 8).permute(0, 1, 3, 2).reshape(T, B8 * 8, 3 * H)[:, :B].float()
            dx = (daf @ kernel.t()).permute(1, 0, 2).contiguous()
        return dx, d_wih.view(1, 3 * H), d_rec, d_bias, None


class _LSTM128v2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, kernel, recurrent, bias, return_sequences):
        B, T, _ = x.shape
        dev = x.device
        x2 = x.reshape(B, T).contiguous()
        need_bwd = any(ctx.needs_input_grad)
        out = torch.empty((B, T, H) if return_sequences else (B, H), dtype=torch.float32, device=dev)
        img = pack_image(recurrent, kernel, bias, 4)
        if need_bwd:
            hs, gates, cs, dz = lstm_v2_buffers(B, T, dev)
        else:
            hs = gates = cs = dz = None
        lstm_v2_forward(x2, img, out, hs, gates, cs, B, T, return_sequences, need_bwd)
        if need_bwd:
            ctx.save_for_backward(x2, kernel, img, hs, gates, cs, dz)
            ctx.return_sequences = return_sequences
            ctx.dims = (B, T)
        return out

    @staticmethod
    def backward(ctx, dout):
        x2, kernel, img, hs, gates, cs, dz = ctx.saved_tensors
        B, T = ctx.dims
        dev = x2.device
        small = torch.zeros(8 * H, dtype=torch.float32, device=dev)
        d_wih, d_bias = small[:4 * H], small[4 * H:]
        d_rec = torch.empty((H, 4 * H), dtype=torch.float32, device=dev)
        lstm_v2_backward(dout.contiguous(), x2, img, gates, cs, dz, d_wih, d_bias, d_rec, B, T, ctx.return_sequences, hs)
        dx = None
        if ctx.needs_input_grad[0]:
            B8 = (B + 7) // 8
            dzf = dz[:T * B8 * 4 * H * 8].view(T, B8, 4 * H, 8).permute(0, 1, 3, 2).reshape(T, B8 * 8, 4 * H)[:, :B].float()
            dx = (dzf @ kernel.t()).permute(1, 0, 2).contiguous()
        return dx, d_wih.view(1, 4 * H), d_rec, d_bias, None


def lstm_layer(x, kernel, recurrent, bias, return_sequences=False):
    return _LSTM128v2.apply(x, kernel, recurrent, bias, bool(return_sequences))


def gru_layer(x, kernel, recurrent, bias, return_sequences=False):
    return _GRU128v2.apply(x, kernel, recurrent, bias, bool(return_sequences))
