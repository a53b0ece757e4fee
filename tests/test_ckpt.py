import numpy as np
import torch

from distributeddeeplearningspark_b200.ckpt import load_checkpoint, load_weights, save_checkpoint, save_weights
from distributeddeeplearningspark_b200.models import mnist_cnn, nyiso_lstm
from distributeddeeplearningspark_b200.optim import KerasAdam
from distributeddeeplearningspark_b200.trainers import StepEngine


def test_keras_layout_npz_names_and_shapes(tmp_path):
    m = mnist_cnn(seed=0)
    p = save_weights(str(tmp_path / 'cnn'), m)
    z = np.load(p)
    assert z['conv2d_1/kernel:0'].shape == (3, 3, 1, 32) and z['dense_1/kernel:0'].shape == (4608, 225)
    l = nyiso_lstm(seed=0)
    z = np.load(save_weights(str(tmp_path / 'lstm.npz'), l))
    assert z['lstm_1/kernel:0'].shape == (1, 512) and z['lstm_1/recurrent_kernel:0'].shape == (128, 512)
    assert np.all(z['lstm_1/bias:0'][128:256] == 1.0)                    # gate order i,f,c,o
    m2 = load_weights(str(tmp_path / 'cnn'))
    for a, b in zip(m.get_weights(), m2.get_weights()):
        np.testing.assert_array_equal(a, b)


def test_resume_continues_bit_identically(tmp_path):
    torch.manual_seed(0)
    x = torch.rand(64, 25, 1); y = torch.rand(64, 1)

    def run(n_steps, engine):
        for i in range(n_steps):
            engine.step(x[(i % 4) * 16:(i % 4 + 1) * 16], y[(i % 4) * 16:(i % 4 + 1) * 16])

    a = StepEngine(nyiso_lstm(units=8, seed=1), 'adam', 'mean_squared_error', device='cpu')
    run(6, a)
    ref = a.model.get_weights()

    b = StepEngine(nyiso_lstm(units=8, seed=1), 'adam', 'mean_squared_error', device='cpu')
    run(3, b)
    save_checkpoint(str(tmp_path / 'ck'), b.model, b.optimizer, step=3, rank=0)
    c = StepEngine(nyiso_lstm(units=8, seed=99), 'adam', 'mean_squared_error', device='cpu')
    _, meta = load_checkpoint(str(tmp_path / 'ck'), c.model, c.optimizer)
    assert meta['step'] == 3 and int(c.optimizer.step_count) == 3
    for i in range(3, 6):
        c.step(x[(i % 4) * 16:(i % 4 + 1) * 16], y[(i % 4) * 16:(i % 4 + 1) * 16])
    for u, v in zip(ref, c.model.get_weights()):
        np.testing.assert_array_equal(u, v)


def test_non_root_rank_does_not_write(tmp_path):
    assert save_checkpoint(str(tmp_path / 'x'), nyiso_lstm(units=4), None, rank=1) is None
    assert not (tmp_path / 'x.npz').exists()
