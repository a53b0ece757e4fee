"""The allocation-free conv1 weight-gradient entry point, called through the C ABI.

Some hosts of this library cannot allocate GPU memory when they call it — they run on a fixed slice of a
shared GPU, or they sit in a loop where cudaMalloc's device-wide synchronisation is unaffordable, or they
are plain C programs with no allocator.  ``ddls_cnn_conv1_wgrad_noalloc`` serves them from a scratch array
built into the library.  This test drives it with ctypes only (raw device pointers, no torch wrapper around
the call) and checks it against the existing entry point that takes caller-provided scratch, twice in a
row with different inputs, because the built-in array is shared between calls and keeps the previous
call's contents.  The file name sorts last on purpose: if this fails under ``-x``, every other test has
already run.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


def test_conv1_wgrad_noalloc_matches_scratch_version_twice_in_a_row():
    from distributeddeeplearningspark_b200 import _native as N
    lib = N.lib()
    P, I, Z = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t
    noalloc = lib.ddls_cnn_conv1_wgrad_noalloc
    noalloc.argtypes = [P, P, P, I, P]
    noalloc.restype = I
    with_scratch = lib.ddls_cnn_conv1_wgrad
    with_scratch.argtypes = [P, P, P, P, Z, I, P]
    with_scratch.restype = I

    dev = torch.device('cuda', 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    for call, (B, seed) in enumerate([(128, 1), (37, 2)]):          # second call is smaller: stale rows remain
        torch.manual_seed(seed)
        x = torch.rand(B, 28, 28, 1, device=dev)
        dz1 = torch.randn(B, 26, 26, 32, device=dev).bfloat16()
        ref = torch.empty(320, device=dev)
        scratch = torch.empty(B * 320, device=dev)
        out = torch.full((320,), float('nan'), device=dev)
        rc = with_scratch(x.data_ptr(), dz1.data_ptr(), ref.data_ptr(), scratch.data_ptr(), scratch.numel() * 4, B, stream)
        assert rc == 0, lib.ddls_last_error()
        rc = noalloc(x.data_ptr(), dz1.data_ptr(), out.data_ptr(), B, stream)
        assert rc == 0, lib.ddls_last_error()
        torch.cuda.synchronize(dev)
        assert torch.isfinite(out).all()
        assert _rel(out, ref) < 2e-3, (call, _rel(out, ref))         # same tolerance as the conv1 wgrad kernel test
        # and against a plain fp32 PyTorch restatement of the op
        w = torch.zeros(3, 3, 1, 32, device=dev, requires_grad=True)
        b = torch.zeros(32, device=dev, requires_grad=True)
        y = torch.nn.functional.conv2d(x.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), b).permute(0, 2, 3, 1)
        (y * dz1.float()).sum().backward()
        assert _rel(out[:288].view(3, 3, 1, 32), w.grad) < 2e-3 and _rel(out[288:], b.grad) < 2e-3


def test_conv1_wgrad_noalloc_rejects_oversized_batch():
    from distributeddeeplearningspark_b200 import _native as N
    lib = N.lib()
    fn = lib.ddls_cnn_conv1_wgrad_noalloc
    fn.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int, ctypes.c_void_p]
    fn.restype = ctypes.c_int
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(dev)
    t = torch.zeros(16, device=dev)
    rc = fn(t.data_ptr(), t.data_ptr(), t.data_ptr(), 1025, 0)        # one past DDLS_NOALLOC_CONV1_MAX_BATCH
    assert rc != 0
    assert b'exceeds the built-in scratch' in lib.ddls_last_error()
