"""Persistent tcgen05 LSTM/GRU kernels and fused heads vs the fp32 PyTorch restatement of the Keras math."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from distributeddeeplearningspark_b200.models import layers as L
from distributeddeeplearningspark_b200.models import nyiso_gru, nyiso_lstm


def _rel(a, b):
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


def _make(cell, seed, dev):
    torch.manual_seed(seed)
    G = 4 if cell == 'lstm' else 3
    k = (torch.rand(1, G * 128, device=dev) - 0.5) * 0.4
    r = torch.empty(128, G * 128)
    L.orthogonal_(r)
    r = r.to(dev)
    b = (torch.rand(G * 128, device=dev) - 0.5) * 0.2
    if cell == 'lstm':
        b[128:256] += 1.0
    return [t.requires_grad_(True) for t in (k, r, b)]


@pytest.mark.parametrize('cell', ['lstm', 'gru'])
@pytest.mark.parametrize('B,T,seq', [(32, 25, False), (7, 5, False), (70, 12, True), (32, 1, False), (300, 9, False)])
def test_recurrent_forward_backward_match_fp32(cell, B, T, seq):
    from distributeddeeplearningspark_b200.ops import recurrent as R
    dev = torch.device('cuda', 0)
    k, r, b = _make(cell, 1, dev)
    x = torch.rand(B, T, 1, device=dev)
    ref_fn = L.lstm_reference if cell == 'lstm' else L.gru_reference
    nat_fn = R.lstm_layer if cell == 'lstm' else R.gru_layer
    ref = ref_fn(x, k, r, b, seq)
    w = torch.randn_like(ref)
    (ref * w).sum().backward()
    gref = [t.grad.clone() for t in (k, r, b)]
    for t in (k, r, b):
        t.grad = None
    out = nat_fn(x, k, r, b, seq)
    assert out.shape == ref.shape
    assert _rel(out, ref.detach()) < 2e-2, _rel(out, ref.detach())
    (out * w).sum().backward()
    for name, g, gr in zip(('kernel', 'recurrent', 'bias'), (k.grad, r.grad, b.grad), gref):
        assert g.shape == gr.shape
        assert _rel(g, gr) < 4e-2, (name, _rel(g, gr))


@pytest.mark.parametrize('tile', [16, 32, 64])
@pytest.mark.parametrize('cell', ['lstm', 'gru'])
def test_recurrent_v2_every_tile_size(cell, tile, monkeypatch):
    from distributeddeeplearningspark_b200.ops import recurrent as R
    monkeypatch.setenv('DDLS_RNN_TILE', str(tile))
    dev = torch.device('cuda', 0)
    k, r, b = _make(cell, 3, dev)
    B, T = 150, 11
    x = torch.rand(B, T, 1, device=dev)
    ref = (L.lstm_reference if cell == 'lstm' else L.gru_reference)(x, k, r, b, False)
    w = torch.randn_like(ref)
    (ref * w).sum().backward()
    gref = [t.grad.clone() for t in (k, r, b)]
    for t in (k, r, b):
        t.grad = None
    out = (R.lstm_layer if cell == 'lstm' else R.gru_layer)(x, k, r, b, False)
    assert _rel(out, ref.detach()) < 2e-2
    (out * w).sum().backward()
    for g, gr in zip((k.grad, r.grad, b.grad), gref):
        assert _rel(g, gr) < 4e-2


def test_dense_mse_head_matches_autograd():
    from distributeddeeplearningspark_b200.ops import heads
    dev = torch.device('cuda', 0)
    torch.manual_seed(0)
    for B, H, O in ((32, 128, 1), (77, 128, 3), (5, 64, 1), (300, 520, 1), (130, 1024, 2)):   # last two: multi-CTA path
        h = torch.randn(B, H, device=dev, requires_grad=True)
        w = (torch.randn(H, O, device=dev) * 0.1).requires_grad_(True)
        b = torch.randn(O, device=dev).requires_grad_(True)
        y = torch.randn(B, O, device=dev)
        ref = (((h @ w + b) - y) ** 2).mean(-1).mean()
        ref.backward()
        gr = [t.grad.clone() for t in (h, w, b)]
        for t in (h, w, b):
            t.grad = None
        loss, pred = heads.dense_mse(h, w, b, y)
        torch.testing.assert_close(loss, ref.detach(), rtol=1e-4, atol=1e-6)
        torch.testing.assert_close(pred, (h @ w + b).detach(), rtol=1e-4, atol=1e-5)
        (3.0 * loss).backward()                      # non-unit seed must be honoured outside the step engine
        for g, r_ in zip((h.grad, w.grad, b.grad), gr):
            torch.testing.assert_close(g, 3.0 * r_, rtol=1e-3, atol=5e-6)
        torch.testing.assert_close(heads.dense_linear(h.detach(), w.detach(), b.detach()), pred)


def test_dense_softmax_xent_head_matches_autograd():
    from distributeddeeplearningspark_b200.ops import heads
    from distributeddeeplearningspark_b200.trainers import categorical_crossentropy
    dev = torch.device('cuda', 0)
    torch.manual_seed(0)
    B, F, C = 128, 225, 10
    a = torch.relu(torch.randn(B, F, device=dev)).requires_grad_(True)
    w = (torch.randn(F, C, device=dev) * 0.1).requires_grad_(True)
    b = torch.randn(C, device=dev).requires_grad_(True)
    y = torch.eye(C, device=dev)[torch.randint(0, C, (B,), device=dev)]
    ref = categorical_crossentropy(torch.softmax(a @ w + b, -1), y)
    ref.backward()
    gr = [t.grad.clone() for t in (a, w, b)]
    for t in (a, w, b):
        t.grad = None
    loss, probs = heads.dense_softmax_xent(a, w, b, y)
    torch.testing.assert_close(loss, ref.detach(), rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(probs, torch.softmax(a @ w + b, -1).detach(), rtol=1e-3, atol=1e-6)
    loss.backward()
    for g, r_ in zip((a.grad, w.grad, b.grad), gr):
        torch.testing.assert_close(g, r_, rtol=2e-3, atol=1e-6)


@pytest.mark.parametrize('make', [nyiso_lstm, nyiso_gru])
def test_step_engine_trains_nyiso_models_with_cuda_graph(make):
    from distributeddeeplearningspark_b200 import _native
    from distributeddeeplearningspark_b200.data import nyiso_windows_arrays
    from distributeddeeplearningspark_b200.trainers import StepEngine
    dev = torch.device('cuda', 0)
    model = make(seed=0).to(dev)
    eng = StepEngine(model, optimizer='adam', loss='mean_squared_error', device=dev, use_cuda_graph=True,
                     optimizer_kwargs={'lr': 5e-3})
    x, y = nyiso_windows_arrays(32 * 40, seed=0)
    x, y = torch.tensor(x, device=dev), torch.tensor(y, device=dev)
    c0 = _native.launch_count()
    losses = []
    for ep in range(3):
        for i in range(40):
            losses.append(eng.step(x[i * 32:(i + 1) * 32], y[i * 32:(i + 1) * 32]).clone())
    losses = torch.stack(losses).cpu().numpy()
    assert eng._graph is not None                       # the step really ran as a replayed CUDA graph
    assert _native.launch_count() > c0
    assert np.isfinite(losses).all() and losses[-20:].mean() < 0.5 * losses[:5].mean()
    # prediction path (fused inference head) agrees with the torch reference equations
    with torch.no_grad():
        pred = model(x[:32])
        h = (L.lstm_reference if make is nyiso_lstm else L.gru_reference)(
            x[:32], model.layers[0].kernel, model.layers[0].recurrent_kernel, model.layers[0].bias)
        ref = h @ model.layers[1].kernel + model.layers[1].bias
    assert float((pred - ref).abs().max()) < 2e-2


@pytest.mark.parametrize('seq', ['1', '0'])
def test_stacked_big_lstm_direct_step_matches_autograd_reference(seq, monkeypatch):
    """Large-hidden stacked LSTM vs torch autograd on the Keras equations: whole-sequence persistent kernels
    (seq=1, csrc/lstm_big_seq.cu) and the per-step kernels (seq=0, csrc/lstm_big_tma.cu)."""
    monkeypatch.setenv('DDLS_LSTM_BIG_SEQ', seq)
    import distributeddeeplearningspark_b200 as ddls
    from distributeddeeplearningspark_b200.ops.fused import DirectStep
    from distributeddeeplearningspark_b200.trainers import mean_squared_error
    dev = torch.device('cuda', 0)
    torch.manual_seed(0)
    B, T = 200, 7            # two batch tiles (one ragged), 4-CTA clusters in BPTT
    m = ddls.models.stacked_lstm(units=256, len_seq=T, num_layers=2, seed=5).to(dev)
    x = torch.rand(B, T, 1, device=dev); y = torch.rand(B, 1, device=dev)
    h = x
    for l in m.layers[:-1]:
        h = L.lstm_reference(h, l.kernel, l.recurrent_kernel, l.bias, l.return_sequences)
    ref = mean_squared_error(h @ m.layers[-1].kernel + m.layers[-1].bias, y)
    ref.backward()
    gref = [p.grad.clone() for p in m.parameters()]
    opt = ddls.optim.KerasAdam(m.parameters())
    assert DirectStep.supported(m, opt, 'mean_squared_error', x)
    ds = DirectStep(m, opt, 'mean_squared_error')
    assert ds.kind == 'stacked_lstm'
    loss = ds.run(x, y)
    torch.cuda.synchronize()
    assert abs(float(loss) - float(ref)) < 2e-2 * max(abs(float(ref)), 1e-3), (float(loss), float(ref))
    for (n, p), g in zip(m.named_parameters(), gref):
        l2 = float((p.grad - g).norm() / (g.norm() + 1e-12))
        assert l2 < 5e-2, (n, l2)
    with torch.no_grad():
        pred = m(x)
    assert float((pred - (h @ m.layers[-1].kernel + m.layers[-1].bias)).abs().max()) < 2e-2


def test_stacked_lstm_at_its_benchmark_size_tracks_fp32_over_168_steps():
    """BASELINE.json config 4 at full size (2 x LSTM(1024), T = 168, batch 512): bf16 recurrent state and bf16 GEMM operands
    against the fp32 restatement of the Keras equations over all 168 dependent steps.  Stated drift tolerance: loss within
    0.1 %, every gradient within 2 % relative L2 (measured on B200: 0.003 % and <= 0.5 %, profiles/stacked_lstm_drift.json)."""
    import json, os
    import distributeddeeplearningspark_b200 as ddls
    from distributeddeeplearningspark_b200.ops.fused import DirectStep
    from distributeddeeplearningspark_b200.trainers import mean_squared_error
    dev = torch.device('cuda', 0)
    torch.manual_seed(0)
    B, T = 512, 168
    m = ddls.models.stacked_lstm(units=1024, len_seq=T, num_layers=2, seed=5).to(dev)
    x = torch.rand(B, T, 1, device=dev); y = torch.rand(B, 1, device=dev)
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False                      # a true fp32 reference
    try:
        h = x
        for l in m.layers[:-1]:
            h = L.lstm_reference(h, l.kernel, l.recurrent_kernel, l.bias, l.return_sequences)
        ref = mean_squared_error(h @ m.layers[-1].kernel + m.layers[-1].bias, y)
        ref.backward()
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    gref = [p.grad.clone() for p in m.parameters()]
    del h
    torch.cuda.empty_cache()
    opt = ddls.optim.KerasAdam(m.parameters())
    assert DirectStep.supported(m, opt, 'mean_squared_error', x)
    loss = DirectStep(m, opt, 'mean_squared_error').run(x, y)
    torch.cuda.synchronize()
    drift = {'loss': float(loss), 'loss_fp32': float(ref), 'loss_rel': abs(float(loss) - float(ref)) / abs(float(ref)),
             'grad_rel_l2': {n: float((p.grad - g).norm() / (g.norm() + 1e-20)) for (n, p), g in zip(m.named_parameters(), gref)}}
    os.makedirs('gpurun_out', exist_ok=True)
    json.dump(drift, open('gpurun_out/stacked_lstm_drift.json', 'w'), indent=1)
    assert drift["loss_rel"] < 1e-3, drift
    assert max(drift["grad_rel_l2"].values()) < 2e-2, drift
