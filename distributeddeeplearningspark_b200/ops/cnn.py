"""Autograd wrapper for the MNIST CNN trunk kernels (``csrc/cnn.cu``, ``csrc/cnn_tma.cu`` + the tcgen05 GEMM)."""
from __future__ import annotations

import torch

from .. import _native as N
from .gemm import gemm_bf16, workspace

P, I, Z = N.c_void_p, N.c_int, N.c_size_t
N.register('ddls_cnn_conv1_fwd', [P, P, P, P, I, P])
N.register('ddls_cnn_conv2_pool_fwd', [P, P, P, P, P, I, P])
N.register('ddls_relu_bwd_bias', [P, P, P, P, I, I, I, P])
N.register('ddls_cnn_unpool_bwd', [P, P, P, P, P, I, P])
N.register('ddls_cnn_conv2_wgrad', [P, P, P, P, Z, I, P])
N.register('ddls_cnn_conv2_dgrad', [P, P, P, P, I, P])
N.register('ddls_cnn_conv1_wgrad', [P, P, P, P, Z, I, P])


def conv2_fwd_name() -> str:
    return 'ddls_cnn_conv2_pool_fwd'


def conv2_wgrad_name() -> str:
    return 'ddls_cnn_conv2_wgrad'


def conv2_dgrad_name() -> str:
    return 'ddls_cnn_conv2_dgrad'


FWD_SPLIT_K = 18


def native_available(plan, x: torch.Tensor) -> bool:
    """The fused trunk covers exactly the reference CNN: 28x28x1 input, two 3x3x32 convs, 2x2 pool, Dense."""
    c1, c2, d1 = plan['conv1'], plan['conv2'], plan['dense1']
    ok = (x.is_cuda and x.dtype == torch.float32 and tuple(x.shape[1:]) == (28, 28, 1)
          and c1.filters == 32 and c2.filters == 32 and c1.kernel_size == (3, 3) and c2.kernel_size == (3, 3)
          and d1.kernel.shape[0] == 4608)
    if ok:
        N.lib()
    return ok


class _CnnTrunk(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k1, b1, k2, b2, kd1, bd1):
        B = x.shape[0]
        dev = x.device
        st = N.stream_ptr(dev)
        x = x.contiguous()
        a1 = torch.empty((B, 26, 26, 32), dtype=torch.bfloat16, device=dev)
        pooled = torch.empty((B, 12, 12, 32), dtype=torch.bfloat16, device=dev)
        pidx = torch.empty((B, 12, 12, 32), dtype=torch.uint8, device=dev)
        N.call('ddls_cnn_conv1_fwd', x.data_ptr(), k1.data_ptr(), b1.data_ptr(), a1.data_ptr(), B, st)
        N.call(conv2_fwd_name(), a1.data_ptr(), k2.data_ptr(), b2.data_ptr(), pooled.data_ptr(),
               pidx.data_ptr(), B, st)
        a3 = gemm_bf16(pooled.view(B, 4608), kd1, bias=bd1, relu=True, b_mn_major=True, split_k=FWD_SPLIT_K)
        if any(ctx.needs_input_grad):
            ctx.save_for_backward(x, k2, kd1, a1, pooled, pidx, a3)
        return a3

    @staticmethod
    def backward(ctx, da3):
        x, k2, kd1, a1, pooled, pidx, a3 = ctx.saved_tensors
        B = x.shape[0]
        dev = x.device
        st = N.stream_ptr(dev)
        F = kd1.shape[1]
        ldz = (F + 7) // 8 * 8
        small = torch.zeros(F + 32, dtype=torch.float32, device=dev)
        dbd1, db2 = small[:F], small[F:]
        dz3 = torch.empty((B, ldz), dtype=torch.bfloat16, device=dev)
        N.call('ddls_relu_bwd_bias', da3.contiguous().data_ptr(), a3.data_ptr(), dz3.data_ptr(), dbd1.data_ptr(), B, F,
               ldz, st)
        flat = pooled.view(B, 4608)
        # dW[f, n] = sum_b flat[b, f] * dz3[b, n]   (both operands MN-major, K = batch)
        dkd1 = gemm_bf16(flat, dz3[:, :F], a_mn_major=True, b_mn_major=True)
        # dflat[b, f] = sum_n dz3[b, n] * W[f, n]   (both K-major)
        dflat = gemm_bf16(dz3[:, :F], kd1, out_dtype=torch.bfloat16)
        dz2 = torch.empty((B, 24, 24, 32), dtype=torch.bfloat16, device=dev)
        N.call('ddls_cnn_unpool_bwd', dflat.data_ptr(), pooled.data_ptr(), pidx.data_ptr(), dz2.data_ptr(),
               db2.data_ptr(), B, st)
        ws = workspace(dev, max(N.lib().ddls_cnn_conv2_wgrad_scratch_bytes(B), B * 320 * 4))
        dk2 = torch.empty((3, 3, 32, 32), dtype=torch.float32, device=dev)
        N.call(conv2_wgrad_name(), a1.data_ptr(), dz2.data_ptr(), dk2.data_ptr(), ws.data_ptr(), ws.numel(), B, st)
        dz1 = torch.empty((B, 26, 26, 32), dtype=torch.bfloat16, device=dev)
        N.call(conv2_dgrad_name(), dz2.data_ptr(), k2.data_ptr(), a1.data_ptr(), dz1.data_ptr(), B, st)
        g1 = torch.empty(320, dtype=torch.float32, device=dev)
        N.call('ddls_cnn_conv1_wgrad', x.data_ptr(), dz1.data_ptr(), g1.data_ptr(), ws.data_ptr(), ws.numel(), B, st)
        return None, g1[:288].view(3, 3, 1, 32), g1[288:], dk2, db2, dkd1, dbd1


def trunk(plan, x: torch.Tensor) -> torch.Tensor:
    c1, c2, d1 = plan['conv1'], plan['conv2'], plan['dense1']
    return _CnnTrunk.apply(x, c1.kernel, c1.bias, c2.kernel, c2.bias, d1.kernel, d1.bias)
