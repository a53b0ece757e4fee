"""GPU numerics: every hand-written kernel against a plain fp32 PyTorch / NumPy-oracle restatement."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def dev():
    return torch.device('cuda', 0)


def test_native_library_is_loaded():
    from distributeddeeplearningspark_b200 import _native
    assert _native.available(), 'libddls_b200.so must be built in-tree'
    assert _native.lib().ddls_version() >= 100


@pytest.mark.parametrize('M,N,K', [(128, 32, 64), (256, 96, 192), (200, 225, 4608), (128, 512, 128), (77, 10, 40)])
@pytest.mark.parametrize('a_mn,b_mn', [(False, False), (True, False), (False, True), (True, True)])
def test_umma_gemm_matches_fp32(dev, M, N, K, a_mn, b_mn):
    from distributeddeeplearningspark_b200.ops.gemm import gemm_bf16
    torch.manual_seed(M + N + K)
    A = torch.randn(M, K, device=dev).bfloat16()
    B = torch.randn(N, K, device=dev).bfloat16()
    ref = A.float() @ B.float().t()
    a = A.t().contiguous() if a_mn else A
    b = B.t().contiguous() if b_mn else B
    out = gemm_bf16(a, b, a_mn_major=a_mn, b_mn_major=b_mn)
    torch.testing.assert_close(out, ref, rtol=2e-3, atol=2e-3 * K ** 0.5)


def test_umma_gemm_epilogue_and_split_k(dev):
    from distributeddeeplearningspark_b200.ops.gemm import gemm_bf16
    torch.manual_seed(1)
    A = torch.randn(128, 4608, device=dev).bfloat16()
    B = (torch.randn(225, 4608, device=dev) * 0.02).bfloat16()
    bias = torch.randn(225, device=dev)
    ref = torch.relu(A.float() @ B.float().t() + bias)
    out = gemm_bf16(A, B, bias=bias, relu=True, out_dtype=torch.bfloat16)
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2)
    out2 = gemm_bf16(A, B, split_k=8)
    torch.testing.assert_close(out2, A.float() @ B.float().t(), rtol=2e-3, atol=5e-3)


@pytest.mark.parametrize('M,N,K', [(256, 64, 64), (300, 264, 200), (1024, 512, 1000), (640, 1024, 4096)])
@pytest.mark.parametrize('a_mn,b_mn', [(False, False), (True, False), (False, True), (True, True)])
def test_tma_gemm_matches_fp32(dev, M, N, K, a_mn, b_mn):
    """The persistent TMA / warp-specialised kernel (csrc/tma_gemm.cu), 128-byte-swizzled operands in all four
    major combinations, ragged M/N/K tails, against an fp32 matmul of the same bf16 inputs."""
    from distributeddeeplearningspark_b200 import _native as NV
    from distributeddeeplearningspark_b200.ops.gemm import gemm_bf16
    torch.manual_seed(M + N + K)
    A = torch.randn(M, K, device=dev).bfloat16()
    B = torch.randn(N, K, device=dev).bfloat16()
    ref = A.float() @ B.float().t()
    a = A.t().contiguous() if a_mn else A
    b = B.t().contiguous() if b_mn else B
    n0 = NV.launch_count()
    out = gemm_bf16(a, b, a_mn_major=a_mn, b_mn_major=b_mn)
    assert NV.launch_count() == n0 + 1
    torch.testing.assert_close(out, ref, rtol=2e-3, atol=2e-3 * K ** 0.5)
    old = gemm_bf16(a, b, a_mn_major=a_mn, b_mn_major=b_mn, no_tma=True)
    torch.testing.assert_close(out, old, rtol=1e-4, atol=1e-3)


def test_tma_gemm_epilogues(dev):
    from distributeddeeplearningspark_b200.ops.gemm import gemm_bf16
    torch.manual_seed(3)
    A = torch.randn(384, 520, device=dev).bfloat16()
    B = (torch.randn(272, 520, device=dev) * 0.05).bfloat16()
    bias = torch.randn(272, device=dev)
    ref = A.float() @ B.float().t()
    out = gemm_bf16(A, B, bias=bias, relu=True, out_dtype=torch.bfloat16)
    torch.testing.assert_close(out.float(), torch.relu(ref + bias), rtol=2e-2, atol=2e-2)
    acc = torch.ones(384, 272, device=dev)
    gemm_bf16(A, B, out=acc, accumulate=True)
    torch.testing.assert_close(acc, ref + 1.0, rtol=2e-3, atol=5e-3)
    wide = torch.zeros(384, 300, device=dev)                     # strided output view
    gemm_bf16(A, B, out=wide[:, :272])
    torch.testing.assert_close(wide[:, :272], ref, rtol=2e-3, atol=5e-3)
    assert float(wide[:, 272:].abs().max()) == 0.0


@pytest.mark.parametrize('kind', ['adam', 'adagrad', 'sgd'])
def test_fused_optimizers_match_oracle(dev, kind):
    from distributeddeeplearningspark_b200.optim import KerasAdagrad, KerasAdam, KerasSGD
    from distributeddeeplearningspark_b200.oracle import keras_math as km
    rng = np.random.default_rng(0)
    shapes = [(33, 7), (129,), (5, 3, 2)]
    p0 = [rng.normal(size=s) for s in shapes]
    params = [torch.nn.Parameter(torch.tensor(p, dtype=torch.float32, device=dev)) for p in p0]
    ref_p = [p.copy() for p in p0]
    opt = {'adam': KerasAdam, 'adagrad': KerasAdagrad, 'sgd': KerasSGD}[kind](params)
    ref = {'adam': km.KerasAdam, 'adagrad': km.KerasAdagrad, 'sgd': km.KerasSGD}[kind]()
    for step in range(7):
        grads = [rng.normal(size=s) for s in shapes]
        opt.zero_grad()
        for p, g in zip(params, grads):
            p.grad.copy_(torch.tensor(g, dtype=torch.float32))
        opt.step()
        ref.step(ref_p, grads)
    assert int(opt.step_count.item()) == 7
    for a, b in zip(ref_p, params):
        np.testing.assert_allclose(a, b.detach().cpu().numpy(), rtol=3e-5, atol=3e-6)


def test_fused_adam_zeroes_grads_and_replays_in_cuda_graph(dev):
    from distributeddeeplearningspark_b200.optim import KerasAdam
    p = torch.nn.Parameter(torch.ones(1000, device=dev))
    opt = KerasAdam([p]); opt.zero_grads_in_step = True
    g_static = torch.full((1000,), 0.5, device=dev)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        p.grad.copy_(g_static); opt.step()
    torch.cuda.current_stream().wait_stream(s)
    assert float(p.grad.abs().max()) == 0.0
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        p.grad.copy_(g_static); opt.step()
    for _ in range(4):
        graph.replay()
    torch.cuda.synchronize()
    assert int(opt.step_count.item()) == 5          # counter lives on the device: replays advance it
    ref = torch.ones(1000, dtype=torch.float64); m = torch.zeros_like(ref); v = torch.zeros_like(ref)
    for t in range(1, 6):
        m = 0.9 * m + 0.1 * 0.5; v = 0.999 * v + 0.001 * 0.25
        ref = ref - 1e-3 * (1 - 0.999 ** t) ** 0.5 / (1 - 0.9 ** t) * m / (v.sqrt() + 1e-7)
    np.testing.assert_allclose(p.detach().cpu().numpy(), ref.numpy(), rtol=1e-5)


def test_mnist_preprocess_kernel(dev):
    from distributeddeeplearningspark_b200.ops.preprocess import mnist_preprocess
    px = torch.randint(0, 256, (64, 784), dtype=torch.uint8)
    lab = torch.randint(0, 10, (64,), dtype=torch.int64)
    x, y = mnist_preprocess(px.to(dev), lab.to(dev))
    xr, yr = mnist_preprocess(px, lab)
    torch.testing.assert_close(x.cpu(), xr); torch.testing.assert_close(y.cpu(), yr)
    assert x.shape == (64, 28, 28, 1) and float(x.max()) <= 1.02 + 1e-6      # o_max = 250 like the reference
