"""Oracle self-consistency: hand-derived NumPy backward vs torch autograd of the same equations."""
import numpy as np
import torch

from distributeddeeplearningspark_b200.models import layers as L
from distributeddeeplearningspark_b200.oracle import keras_math as km

torch.manual_seed(0)
RNG = np.random.default_rng(0)


def _t(a, grad=False):
    t = torch.tensor(a, dtype=torch.float64)
    t.requires_grad_(grad)
    return t


def test_hard_sigmoid_matches_keras_definition():
    x = np.array([-3.0, -2.5, -1.0, 0.0, 1.0, 2.5, 3.0])
    np.testing.assert_allclose(km.hard_sigmoid(x), [0, 0, 0.3, 0.5, 0.7, 1, 1])


def test_lstm_backward_matches_autograd():
    B, T, D, H = 3, 5, 1, 8
    k, r, b = RNG.normal(size=(D, 4 * H)), RNG.normal(size=(H, 4 * H)) * 0.5, RNG.normal(size=4 * H) * 0.1
    x = RNG.normal(size=(B, T, D))
    dout = RNG.normal(size=(B, H))
    out, cache = km.lstm_forward(k, r, b, x)
    dx, dk, dr, db = km.lstm_backward(k, r, b, cache, dout)
    tk, tr, tb, tx = _t(k, True), _t(r, True), _t(b, True), _t(x, True)
    o = L.lstm_reference(tx, tk, tr, tb)
    np.testing.assert_allclose(o.detach().numpy(), out, atol=1e-12)
    (o * _t(dout)).sum().backward()
    for a, g in ((dk, tk.grad), (dr, tr.grad), (db, tb.grad), (dx, tx.grad)):
        np.testing.assert_allclose(a, g.numpy(), atol=1e-10)


def test_lstm_sequences_backward_matches_autograd():
    B, T, D, H = 2, 4, 3, 6
    k, r, b = RNG.normal(size=(D, 4 * H)), RNG.normal(size=(H, 4 * H)) * 0.5, RNG.normal(size=4 * H) * 0.1
    x = RNG.normal(size=(B, T, D))
    dout = RNG.normal(size=(B, T, H))
    out, cache = km.lstm_forward(k, r, b, x, return_sequences=True)
    dx, dk, dr, db = km.lstm_backward(k, r, b, cache, dout)
    tk, tr, tb, tx = _t(k, True), _t(r, True), _t(b, True), _t(x, True)
    o = L.lstm_reference(tx, tk, tr, tb, return_sequences=True)
    np.testing.assert_allclose(o.detach().numpy(), out, atol=1e-12)
    (o * _t(dout)).sum().backward()
    for a, g in ((dk, tk.grad), (dr, tr.grad), (db, tb.grad), (dx, tx.grad)):
        np.testing.assert_allclose(a, g.numpy(), atol=1e-10)


def test_gru_backward_matches_autograd():
    B, T, D, H = 3, 6, 1, 8
    k, r, b = RNG.normal(size=(D, 3 * H)), RNG.normal(size=(H, 3 * H)) * 0.5, RNG.normal(size=3 * H) * 0.1
    x = RNG.normal(size=(B, T, D))
    dout = RNG.normal(size=(B, H))
    out, cache = km.gru_forward(k, r, b, x)
    dx, dk, dr, db = km.gru_backward(k, r, b, cache, dout)
    tk, tr, tb, tx = _t(k, True), _t(r, True), _t(b, True), _t(x, True)
    o = L.gru_reference(tx, tk, tr, tb)
    np.testing.assert_allclose(o.detach().numpy(), out, atol=1e-12)
    (o * _t(dout)).sum().backward()
    for a, g in ((dk, tk.grad), (dr, tr.grad), (db, tb.grad), (dx, tx.grad)):
        np.testing.assert_allclose(a, g.numpy(), atol=1e-10)


def test_mnist_cnn_grads_match_autograd():
    from distributeddeeplearningspark_b200.models import mnist_cnn
    from distributeddeeplearningspark_b200.trainers import categorical_crossentropy
    m = mnist_cnn(seed=1).double()
    B = 4
    x = RNG.uniform(0, 1, size=(B, 28, 28, 1))
    labels = RNG.integers(0, 10, size=B)
    y = np.eye(10)[labels]
    w = [a.astype(np.float64) for a in m.get_weights()]
    # non-zero biases so their gradients are exercised
    for i in (1, 3, 5, 7):
        w[i] = RNG.normal(size=w[i].shape) * 0.05
    m.set_weights(w)
    loss, grads, probs = km.mnist_cnn_loss_and_grads(w, x, y)
    out = m(_t(x))
    np.testing.assert_allclose(out.detach().numpy(), probs, atol=1e-10)
    tl = categorical_crossentropy(out, _t(y))
    assert abs(float(tl) - loss) < 1e-10
    tl.backward()
    for g, (_, p) in zip(grads, m.named_keras_weights()):
        np.testing.assert_allclose(g, p.grad.numpy(), atol=1e-9)


def test_rnn_regressor_grads_match_autograd():
    from distributeddeeplearningspark_b200.models import nyiso_gru, nyiso_lstm
    from distributeddeeplearningspark_b200.trainers import mean_squared_error
    for cell, make in (('lstm', nyiso_lstm), ('gru', nyiso_gru)):
        m = make(units=16, len_seq=7, seed=2).double()
        w = [a.astype(np.float64) for a in m.get_weights()]
        x = RNG.uniform(0, 1, size=(5, 7, 1)); y = RNG.uniform(0, 1, size=(5, 1))
        loss, grads, pred = km.rnn_regressor_loss_and_grads(cell, w, x, y)
        out = m(_t(x))
        np.testing.assert_allclose(out.detach().numpy(), pred, atol=1e-12)
        tl = mean_squared_error(out, _t(y))
        assert abs(float(tl) - loss) < 1e-12
        tl.backward()
        for g, (_, p) in zip(grads, m.named_keras_weights()):
            np.testing.assert_allclose(g, p.grad.numpy(), atol=1e-10)


def test_keras_adam_and_adagrad_match_flat_optimizers():
    from distributeddeeplearningspark_b200.optim import KerasAdam, KerasAdagrad
    for ocls, tcls in ((km.KerasAdam, KerasAdam), (km.KerasAdagrad, KerasAdagrad)):
        p0 = [RNG.normal(size=(5, 3)), RNG.normal(size=(7,))]
        params_np = [p.copy() for p in p0]
        params_t = [torch.nn.Parameter(torch.tensor(p, dtype=torch.float32)) for p in p0]
        o_np, o_t = ocls(), tcls(params_t)
        for step in range(5):
            grads = [RNG.normal(size=p.shape) for p in p0]
            o_np.step(params_np, grads)
            o_t.zero_grad()
            for p, g in zip(params_t, grads):
                p.grad.copy_(torch.tensor(g, dtype=torch.float32))
            o_t.step()
        for a, b in zip(params_np, params_t):
            np.testing.assert_allclose(a, b.detach().numpy(), rtol=2e-5, atol=2e-6)
