"""tcgen05 GEMM wrapper (``csrc/umma_gemm.cu``)."""
from __future__ import annotations

from typing import Optional

import torch

from .. import _native as N

RELU, BIAS, OUT_BF16 = 1, 2, 4


def gemm_bf16(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False,
              out_dtype=torch.float32, a_mn_major: bool = False, b_mn_major: bool = False, split_k: int = 1,
              out: Optional[torch.Tensor] = None, debug: int = 0) -> torch.Tensor:
    """C[M,N] = A·Bᵀ.  ``a`` is [M,K] (or [K,M] if ``a_mn_major``); ``b`` is [N,K] (or [K,N] if ``b_mn_major``)."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_cuda and b.is_cuda
    assert a.stride(-1) == 1 and b.stride(-1) == 1
    M, K = (a.shape[1], a.shape[0]) if a_mn_major else (a.shape[0], a.shape[1])
    Nn, Kb = (b.shape[1], b.shape[0]) if b_mn_major else (b.shape[0], b.shape[1])
    assert K == Kb, (a.shape, b.shape)
    if out is None:
        out = (torch.zeros if split_k > 1 else torch.empty)((M, Nn), dtype=out_dtype, device=a.device)
    flags = (RELU if relu else 0) | (BIAS if bias is not None else 0) | (OUT_BF16 if out.dtype == torch.bfloat16 else 0)
    N.call('ddls_gemm_bf16', a.data_ptr(), b.data_ptr(), out.data_ptr(), N.ptr(bias), M, Nn, K, a.stride(0),
           b.stride(0), out.stride(0), int(a_mn_major), int(b_mn_major), split_k, flags, debug,
           N.stream_ptr(a.device))
    return out
