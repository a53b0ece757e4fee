"""tcgen05 GEMM wrapper (``csrc/umma_gemm.cu``)."""
from __future__ import annotations

from typing import Dict, Optional

import torch

from .. import _native as N

RELU, BIAS, OUT_BF16, A_F32, B_F32, ACCUMULATE, FORCE_V1, NO_TMA = 1, 2, 4, 16, 32, 128, 256, 512

_WORKSPACES: Dict[torch.device, torch.Tensor] = {}


def workspace(device, nbytes: int) -> torch.Tensor:
    """Grow-only per-device scratch buffer for split-K partial tiles (stable address once warm, so it is
    safe inside CUDA graphs after the warm-up steps)."""
    ws = _WORKSPACES.get(device)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 8 << 20), dtype=torch.uint8, device=device)
        _WORKSPACES[device] = ws
    return ws


def gemm_bf16(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False,
              out_dtype=torch.float32, a_mn_major: bool = False, b_mn_major: bool = False, split_k: int = 1,
              out: Optional[torch.Tensor] = None, accumulate: bool = False, no_tma: bool = False) -> torch.Tensor:
    """C[M,N] = A·Bᵀ.  ``a`` is [M,K] (or [K,M] if ``a_mn_major``); ``b`` is [N,K] (or [K,N] if ``b_mn_major``).
    Operands may be bf16 or fp32 (fp32 is rounded to bf16 while staging into shared memory)."""
    assert a.is_cuda and b.is_cuda and a.stride(-1) == 1 and b.stride(-1) == 1
    assert a.dtype in (torch.bfloat16, torch.float32) and b.dtype in (torch.bfloat16, torch.float32)
    M, K = (a.shape[1], a.shape[0]) if a_mn_major else (a.shape[0], a.shape[1])
    Nn, Kb = (b.shape[1], b.shape[0]) if b_mn_major else (b.shape[0], b.shape[1])
    assert K == Kb, (a.shape, b.shape)
    if out is None:
        out = torch.empty((M, Nn), dtype=out_dtype, device=a.device)
    flags = (RELU if relu else 0) | (BIAS if bias is not None else 0) | (OUT_BF16 if out.dtype == torch.bfloat16 else 0)
    flags |= (A_F32 if a.dtype == torch.float32 else 0) | (B_F32 if b.dtype == torch.float32 else 0)
    flags |= ACCUMULATE if accumulate else 0
    flags |= NO_TMA if no_tma else 0
    split_k = max(1, min(split_k, (K + 63) // 64))
    ws_ptr, ws_bytes = 0, 0
    if split_k > 1:
        ws = workspace(a.device, split_k * M * Nn * 4)
        ws_ptr, ws_bytes = ws.data_ptr(), ws.numel()
    N.call('ddls_gemm_bf16', a.data_ptr(), b.data_ptr(), out.data_ptr(), N.ptr(bias), M, Nn, K, a.stride(0),
           b.stride(0), out.stride(0), int(a_mn_major), int(b_mn_major), split_k, flags, ws_ptr, ws_bytes,
           N.stream_ptr(a.device))
    return out
