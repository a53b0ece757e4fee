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


class _RoundGrad(torch.autograd.Function):
    """Identity whose gradient is rounded to bf16 (the kernels store activation gradients in bf16)."""

    @staticmethod
    def forward(ctx, x):
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return g.bfloat16().float()


def _bf(t):
    return t.bfloat16().float()


def _reference_trunk(m, x):
    """fp32 PyTorch trunk with bf16 rounding at exactly the points where the kernels store bf16
    (a1, conv2/dense weights as tensor-core operands, pooled, and the activation gradients), so ReLU
    masks and pooling arg-max decisions are taken on the same values as in the kernels."""
    import torch.nn.functional as F
    c1, c2, d1 = m.layers[0], m.layers[2], m.layers[6]
    z1 = F.conv2d(x.permute(0, 3, 1, 2), c1.kernel.permute(3, 2, 0, 1), c1.bias)
    a1 = _bf(torch.relu(_RoundGrad.apply(z1)))
    z2 = F.conv2d(a1, _bf(c2.kernel).permute(3, 2, 0, 1), c2.bias)
    p = _bf(torch.relu(F.max_pool2d(z2, 2)))
    flat = _RoundGrad.apply(p.permute(0, 2, 3, 1).reshape(x.shape[0], -1))
    z3 = flat @ _bf(d1.kernel) + d1.bias
    return torch.relu(_RoundGrad.apply(z3))


@pytest.mark.parametrize('B', [128, 5, 37])
def test_cnn_trunk_forward_backward_match_fp32(B):
    from distributeddeeplearningspark_b200.ops import cnn, fused
    dev = torch.device('cuda', 0)
    m = _model(dev)
    torch.manual_seed(B)
    x = torch.rand(B, 28, 28, 1, device=dev)
    a = _reference_trunk(m, x)
    w = torch.randn_like(a)
    (a * w).sum().backward()
    ref_grads = [p.grad.clone() for _, p in m.named_keras_weights()[:6]]
    for _, p in m.named_keras_weights():
        p.grad = None
    # sanity: the rounding-aware reference is itself close to the plain fp32 layers
    plain = x
    for l in m.layers[:8]:
        plain = l(plain)
    assert _rel(a.detach(), plain.detach()) < 2e-2
    plan = fused._plan(m)
    assert plan['kind'] == 'cnn' and cnn.native_available(plan, x)
    out = cnn.trunk(plan, x)
    assert out.shape == a.shape
    assert _rel(out, a.detach()) < 2e-3, _rel(out, a.detach())
    (out * w).sum().backward()
    names = ['conv1.k', 'conv1.b', 'conv2.k', 'conv2.b', 'dense1.k', 'dense1.b']
    grads = [p.grad for _, p in m.named_keras_weights()][:6]
    for n, g, r in zip(names, grads, ref_grads):
        assert g is not None and g.shape == r.shape, n
        l2 = float((g - r).norm() / (r.norm() + 1e-12))
        assert l2 < 1e-2, (n, l2)


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


@pytest.mark.parametrize('name', ['mnist_cnn', 'nyiso_lstm', 'nyiso_gru'])
def test_direct_step_matches_autograd_path(name, monkeypatch):
    """The autograd-free fused step and the autograd Functions are the same math: same losses and weights."""
    import distributeddeeplearningspark_b200 as ddls
    from distributeddeeplearningspark_b200.trainers import StepEngine
    dev = torch.device('cuda', 0)
    torch.manual_seed(0)
    if name == 'mnist_cnn':
        xs = torch.rand(6, 64, 28, 28, 1, device=dev)
        ys = torch.eye(10, device=dev)[torch.randint(0, 10, (6, 64), device=dev)]
    else:
        xs = torch.rand(6, 32, 25, 1, device=dev); ys = torch.rand(6, 32, 1, device=dev)
    out = {}
    for direct in ('1', '0'):
        monkeypatch.setenv('DDLS_DIRECT_STEP', direct)
        m = getattr(ddls.models, name)(seed=3).to(dev)
        eng = StepEngine(m, optimizer=m.optimizer_name, loss=m.loss_name, device=dev, use_cuda_graph=False)
        losses = [float(eng.step(xs[i], ys[i])) for i in range(6)]
        assert (eng._direct is not None) == (direct == '1')
        out[direct] = (losses, torch.cat([p.detach().reshape(-1) for p in m.parameters()]))
        if direct == '1' and name == 'mnist_cnn':      # bf16 shadow of the Dense kernel tracks the fp32 master
            d1 = m.layers[6]
            sh = eng._inner._shadow_by_param[id(d1.kernel)]
            assert torch.equal(sh[:, :225], d1.kernel.detach().bfloat16()) and float(sh[:, 225:].float().abs().max()) == 0
    la, lb = out['1'][0], out['0'][0]
    assert all(abs(a - b) <= 2e-2 * max(abs(b), 1e-3) for a, b in zip(la, lb)), (la, lb)
    wa, wb = out['1'][1], out['0'][1]
    assert float((wa - wb).norm() / wb.norm()) < 2e-3


def test_step_engine_device_dataset_matches_explicit_batches():
    """attach_dataset/step_dataset (in-graph gather with a device-side cursor) == step(x, y) on the same minibatches,
    including the wrap-around at the end of the data and a shuffled order."""
    import distributeddeeplearningspark_b200 as ddls
    from distributeddeeplearningspark_b200.trainers import StepEngine
    dev = torch.device('cuda', 0)
    torch.manual_seed(0)
    Nrows, B, T = 100, 32, 25
    X = torch.rand(Nrows, T, 1, device=dev); Y = torch.rand(Nrows, 1, device=dev)
    perm = torch.randperm(Nrows, device=dev).int()
    ma = ddls.models.nyiso_lstm(seed=3).to(dev); mb = ddls.models.nyiso_lstm(seed=3).to(dev)
    # SGD: linear in the gradient, so the run-to-run rounding noise of the recurrent kernels' atomics is not amplified
    # the way Adam's m/(sqrt(v)+eps) amplifies it
    ea = StepEngine(ma, optimizer='sgd', loss='mean_squared_error', device=dev, use_cuda_graph=True)
    eb = StepEngine(mb, optimizer='sgd', loss='mean_squared_error', device=dev, use_cuda_graph=True)
    ea.attach_dataset(X, Y, B, perm=perm)
    xb, yb = eb.input_buffers(X[:B], Y[:B])
    for k in range(9):                                   # 3 eager warm-ups, the capture, replays; wraps after step 3
        la = ea.step_dataset()
        rows = perm[(torch.arange(B, device=dev) + k * B) % Nrows].long()
        xb.copy_(X[rows]); yb.copy_(Y[rows])
        lb = eb.step(xb, yb)                             # static buffers passed back: no device-to-device copy
        torch.cuda.synchronize()
        assert abs(float(la) - float(lb)) <= 1e-4 * max(1.0, abs(float(lb))), (k, float(la), float(lb))
    assert ea.dataset_cursor() == 9
    for pa, pb in zip(ma.parameters(), mb.parameters()):
        torch.testing.assert_close(pa, pb, rtol=1e-4, atol=1e-5)


def test_step_engine_run_stream_matches_step_by_step():
    """Pipelined host feed (two input slots / graphs, copy stream, losses read one step behind) == step(x, y)."""
    import distributeddeeplearningspark_b200 as ddls
    from distributeddeeplearningspark_b200.trainers import StepEngine
    dev = torch.device('cuda', 0)
    torch.manual_seed(1)
    n, B, T = 11, 32, 25
    hx = [torch.rand(B, T, 1).pin_memory() for _ in range(n)]
    hy = [torch.rand(B, 1).pin_memory() for _ in range(n)]
    ma = ddls.models.nyiso_gru(seed=4).to(dev); mb = ddls.models.nyiso_gru(seed=4).to(dev)
    ea = StepEngine(ma, optimizer='sgd', loss='mean_squared_error', device=dev, use_cuda_graph=True)
    eb = StepEngine(mb, optimizer='sgd', loss='mean_squared_error', device=dev, use_cuda_graph=True)
    la = ea.run_stream(zip(hx, hy), hx[0], hy[0])
    lb = []
    for x, y in zip(hx, hy):
        lb.append(float(eb.step(x.to(dev), y.to(dev))))
    torch.cuda.synchronize()
    assert len(la) == n and all(s['graph'] is not None for s in ea._slots)
    for k, (a, b) in enumerate(zip(la, lb)):
        assert abs(a - b) <= 1e-4 * max(1.0, abs(b)), (k, a, b)
    for pa, pb in zip(ma.parameters(), mb.parameters()):
        torch.testing.assert_close(pa, pb, rtol=1e-4, atol=1e-5)
