"""ctypes binding to ``_lib/libddls_b200.so`` (the hand-written sm_100a kernels + peer-memory runtime).

The library is built in-tree by ``build.py`` (``nvcc -gencode arch=compute_100a,code=sm_100a``) so it
travels with the repo snapshot.  Ops on CUDA tensors fail loudly if it is missing — there is no
silent PyTorch fallback on a GPU box.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_longlong, c_size_t, c_uint, c_ulonglong, c_void_p
from typing import Optional

_LIB: Optional[ctypes.CDLL] = None
LIB_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), '_lib')
LIB_PATH = os.path.join(LIB_DIR, 'libddls_b200.so')


class NativeError(RuntimeError):
    pass


def available() -> bool:
    return os.path.exists(LIB_PATH)


def _proto(lib, name, argtypes, restype=c_int):
    fn = getattr(lib, name)
    fn.argtypes = argtypes
    fn.restype = restype
    return fn


def lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise NativeError(f'{LIB_PATH} is missing: run `python -m distributeddeeplearningspark_b200.build` '
                          f'(nvcc, sm_100a) before using CUDA tensors')
    L = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
    _proto(L, 'ddls_last_error', [], c_char_p)
    _proto(L, 'ddls_version', [], c_int)
    _proto(L, 'ddls_launch_count', [], c_ulonglong)
    if hasattr(L, 'ddls_cnn_conv2_wgrad_scratch_bytes'):
        _proto(L, 'ddls_cnn_conv2_wgrad_scratch_bytes', [c_int], c_size_t)
    for name, args in _PROTOTYPES.items():
        if hasattr(L, name):
            _proto(L, name, args)
    _LIB = L
    return L


P = c_void_p
_PROTOTYPES = {
    'ddls_gemm_bf16': [P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_uint, P, c_size_t, P],
    'ddls_adam_step': [P, P, P, P, P, c_size_t, c_float, c_float, c_float, c_float, c_float, c_int, P],
    'ddls_adagrad_step': [P, P, P, P, c_size_t, c_float, c_float, c_float, c_int, P],
    'ddls_sgd_step': [P, P, P, c_size_t, c_float, c_float, c_int, P],
    'ddls_mnist_preprocess': [P, P, P, P, c_int, c_int, c_int, c_float, c_float, c_float, P],
}


def register(name: str, argtypes) -> None:
    """Let op modules declare prototypes next to their wrappers."""
    _PROTOTYPES[name] = argtypes
    if _LIB is not None and hasattr(_LIB, name):
        _proto(_LIB, name, argtypes)


def check(rc: int, what: str = '') -> None:
    if rc != 0:
        msg = lib().ddls_last_error()
        raise NativeError(f'{what} failed (code {rc}): {msg.decode() if msg else "unknown error"}')


def call(name: str, *args) -> None:
    check(getattr(lib(), name)(*args), name)


def launch_count() -> int:
    return int(lib().ddls_launch_count()) if _LIB is not None or available() else 0


def stream_ptr(device=None) -> int:
    import torch
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t) -> int:
    return 0 if t is None else t.data_ptr()
