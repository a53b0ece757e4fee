"""MNIST CNN trunk kernels vs the fp32 PyTorch restatement of the Keras layers."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from distributeddeeplearningspark_b200.models import mnist_cnn
from distributeddeeplearningspark_b200.trainers import categorical_crossentropy


def _rel(a, b):
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


def _model(dev, seed=0):
    m = mnist_cnn(seed=seed).to(dev)
    with torch.no_grad():       # non-zero biases so their gradients are exercised
        for l in m.layers:
            if hasattr(l, 'bias') and l.bias is not None:
                l.bias.uniform_(-0.05, 0.05)
    return m


@pytest.mark.parametrize('B', [128, 5, 37])
def test_cnn_trunk_forward_backward_match_fp32(B):
    from distributeddeeplearningspark_b200.ops import cnn, fused
    dev = torch.device('cuda', 0)
    m = _model(dev)
    torch.manual_seed(B)
    x = torch.rand(B, 28, 28, 1, device=dev)
    ls = m.layers
    a = x
    for l in ls[:8]:
        a = l(a)                               # torch reference trunk (fp32)
    w = torch.randn_like(a)
    (a * w).sum().backward()
    ref_grads = [p.grad.clone() for _, p in m.named_keras_weights()][:6]
    for _, p in m.named_keras_weights():
        p.grad = None
    plan = fused._plan(m)
    assert plan['kind'] == 'cnn' and cnn.native_available(plan, x)
    out = cnn.trunk(plan, x)
    assert out.shape == a.shape
    assert _rel(out, a.detach()) < 2e-2, _rel(out, a.detach())
    (out * w).sum().backward()
    names = ['conv1.k', 'conv1.b', 'conv2.k', 'conv2.b', 'dense1.k', 'dense1.b']
    grads = [p.grad for _, p in m.named_keras_weights()][:6]
    for n, g, r in zip(names, grads, ref_grads):
        assert g is not None and g.shape == r.shape, n
        assert _rel(g, r) < 5e-2, (n, _rel(g, r))


def test_cnn_loss_and_probs_match_reference_path():
    from distributeddeeplearningspark_b200.ops import fused
    dev = torch.device('cuda', 0)
    m = _model(dev, 1)
    x = torch.rand(64, 28, 28, 1, device=dev)
    y = torch.eye(10, device=dev)[torch.randint(0, 10, (64,), device=dev)]
    a = x
    for l in m.layers:
        a = l(a)
    ref = categorical_crossentropy(a, y)
    loss = fused.try_loss(m, x, y, 'categorical_crossentropy')
    assert abs(float(loss) - float(ref)) < 2e-2 * max(1.0, abs(float(ref)))
    with torch.no_grad():
        probs = m(x)                            # fused inference path
    assert float((probs - a.detach()).abs().max()) < 2e-2
    torch.testing.assert_close(probs.sum(-1), torch.ones(64, device=dev), rtol=1e-4, atol=1e-4)


def test_step_engine_trains_mnist_cnn_with_cuda_graph():
    from distributeddeeplearningspark_b200 import _native
    from distributeddeeplearningspark_b200.data import mnist_like_arrays
    from distributeddeeplearningspark_b200.ops.preprocess import mnist_preprocess
    from distributeddeeplearningspark_b200.trainers import StepEngine
    dev = torch.device('cuda', 0)
    m = mnist_cnn(seed=0).to(dev)
    eng = StepEngine(m, optimizer='adam', loss='categorical_crossentropy', device=dev, use_cuda_graph=True)
    px, lab = mnist_like_arrays(128 * 20, seed=0)
    x, y = mnist_preprocess(torch.from_numpy(px).to(dev), torch.from_numpy(lab).to(dev))
    c0 = _native.launch_count()
    losses = []
    for ep in range(2):
        for i in range(20):
            losses.append(eng.step(x[i * 128:(i + 1) * 128], y[i * 128:(i + 1) * 128]).clone())
    losses = torch.stack(losses).cpu().numpy()
    assert eng._graph is not None and _native.launch_count() > c0
    assert np.isfinite(losses).all() and losses[-5:].mean() < 0.5 * losses[:3].mean()
    with torch.no_grad():
        acc = float((m(x[:512]).argmax(-1) == torch.from_numpy(lab[:512]).to(dev)).float().mean())
    assert acc > 0.8
