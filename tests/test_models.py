import json

import numpy as np
import torch

from distributeddeeplearningspark_b200.models import (Sequential, clone_model, deserialize_model, mnist_cnn,
                                                       nyiso_gru, nyiso_lstm, serialize_model, stacked_lstm)


def test_param_counts_match_reference_summaries():
    # ddl_nyiso_hdi.ipynb:549-553 (50,049), :772-776 (66,689); MNIST CNN derived in SURVEY.md §0.4
    assert nyiso_gru().count_params() == 50049
    assert nyiso_lstm().count_params() == 66689
    assert mnist_cnn().count_params() == 1048853
    assert stacked_lstm().count_params() == 12596225


def test_keras_weight_order_and_shapes():
    m = mnist_cnn()
    names = [n for n, _ in m.named_keras_weights()]
    assert names == ['conv2d_1/kernel:0', 'conv2d_1/bias:0', 'conv2d_2/kernel:0', 'conv2d_2/bias:0',
                     'dense_1/kernel:0', 'dense_1/bias:0', 'dense_2/kernel:0', 'dense_2/bias:0']
    shapes = [w.shape for w in m.get_weights()]
    assert shapes == [(3, 3, 1, 32), (32,), (3, 3, 32, 32), (32,), (4608, 225), (225,), (225, 10), (10,)]
    l = nyiso_lstm()
    assert [w.shape for w in l.get_weights()] == [(1, 512), (128, 512), (512,), (128, 1), (1,)]
    g = nyiso_gru()
    assert [w.shape for w in g.get_weights()] == [(1, 384), (128, 384), (384,), (128, 1), (1,)]


def test_keras_default_initialisers():
    l = nyiso_lstm(seed=0)
    k, r, b = l.get_weights()[:3]
    assert np.all(b[:128] == 0) and np.all(b[128:256] == 1) and np.all(b[256:] == 0)   # unit forget bias
    np.testing.assert_allclose(r @ r.T, np.eye(128), atol=1e-4)                           # orthogonal rows
    limit = np.sqrt(6.0 / (1 + 512))
    assert np.abs(k).max() <= limit + 1e-6                                                # glorot uniform
    m = mnist_cnn(seed=0)
    kd = m.get_weights()[4]
    assert np.abs(kd).max() <= np.sqrt(6.0 / (4608 + 225)) + 1e-6


def test_summary_prints_totals():
    text = nyiso_gru().summary(print_fn=None)
    assert 'Total params: 50,049' in text and 'gru_1 (GRU)' in text and '(None, 128)' in text


def test_serialize_roundtrip_and_json():
    for make in (mnist_cnn, nyiso_lstm, nyiso_gru):
        m = make(seed=3)
        blob = serialize_model(m)
        json.loads(blob['model'])
        m2 = deserialize_model(blob)
        for a, b in zip(m.get_weights(), m2.get_weights()):
            np.testing.assert_array_equal(a, b)
        x = torch.rand((2,) + m.input_shape)
        torch.testing.assert_close(m(x), m2(x))
        assert m2.loss_name == m.loss_name
    assert clone_model(nyiso_gru()).count_params() == 50049


def test_cnn_output_is_softmax_and_flatten_order_hwc():
    m = mnist_cnn(seed=0)
    out = m(torch.rand(3, 28, 28, 1))
    assert out.shape == (3, 10)
    torch.testing.assert_close(out.sum(-1), torch.ones(3))


def test_keras_sgd_momentum_matches_torch_sgd():
    """KerasSGD with (Nesterov) momentum == torch.optim.SGD for a constant learning rate."""
    import torch
    import distributeddeeplearningspark_b200 as ddls
    for nesterov in (False, True):
        torch.manual_seed(0)
        p1 = [torch.nn.Parameter(torch.randn(10)), torch.nn.Parameter(torch.randn(3, 4))]
        p2 = [torch.nn.Parameter(p.detach().clone()) for p in p1]
        o1 = ddls.optim.KerasSGD(p1, lr=0.05, momentum=0.9, nesterov=nesterov)
        o2 = torch.optim.SGD(p2, lr=0.05, momentum=0.9, nesterov=nesterov)
        for _ in range(5):
            gs = [torch.randn_like(p) for p in p1]
            for p, g in zip(p1, gs):
                p.grad.copy_(g)
            for p, g in zip(p2, gs):
                p.grad = g.clone()
            o1.step(); o2.step()
        for a, b in zip(p1, p2):
            torch.testing.assert_close(a.detach(), b.detach(), rtol=1e-5, atol=1e-6)
