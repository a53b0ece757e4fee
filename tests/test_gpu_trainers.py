"""Trainer facades and checkpoints on a CUDA device (the CPU tiers run the same code with DDLS_FORCE_CPU=1)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _nyiso_frames(n_hours=700):
    from distributeddeeplearningspark_b200.data import MinMaxTransformer, make_nyiso_windows, nyiso_like_frame, split_train_test
    df = nyiso_like_frame(n_hours, seed=1)
    lo, hi = df.agg_min('HourAvgLoad'), df.agg_max('HourAvgLoad')
    df = MinMaxTransformer(0, 1, df.agg_min('temperature'), df.agg_max('temperature'), 'temperature', 'NormTemp',
                           is_vector=False).transform(df)
    df = MinMaxTransformer(0, 1, lo, hi, 'HourAvgLoad', 'NormLoad', is_vector=False).transform(df)
    return split_train_test(make_nyiso_windows(df), 120)


def test_trainer_history_on_cuda_is_one_value_per_step():
    """StepEngine returns its static loss tensor on every CUDA-graph replay; the worker must snapshot it per step
    (the history used to read [warm-up losses, last, last, last, ...])."""
    from distributeddeeplearningspark_b200.models import nyiso_lstm
    from distributeddeeplearningspark_b200.trainers import SingleTrainer
    tr, _ = _nyiso_frames()
    t = SingleTrainer(keras_model=nyiso_lstm(seed=0), worker_optimizer='adam', loss='mean_squared_error', num_epoch=4,
                      batch_size=32, features_col='feature_matrix', label_col='labels', optimizer_kwargs={'lr': 0.01})
    t.train(tr)
    h = t.get_history()[0]
    assert len(h) == (tr.count() * 4) // 32
    tail = h[8:]                                        # past the eager warm-up steps: these came from graph replays
    assert len(set(np.round(tail, 9).tolist())) > len(tail) // 2, 'history repeats one aliased loss value'
    assert np.mean(h[-10:]) < 0.7 * np.mean(h[:10])


def test_cnn_checkpoint_resume_on_cuda_refreshes_bf16_shadows(tmp_path):
    """load_checkpoint into a model whose DirectStep (and bf16 shadow of dense_1/kernel) already exists: the next step must
    use the loaded weights (set_weights does not bump the flat buffer's version counter)."""
    from distributeddeeplearningspark_b200.ckpt import load_checkpoint, save_checkpoint
    from distributeddeeplearningspark_b200.models import mnist_cnn
    from distributeddeeplearningspark_b200.trainers import StepEngine
    dev = torch.device('cuda', 0)
    torch.manual_seed(0)
    x = torch.rand(64, 28, 28, 1, device=dev)
    y = torch.nn.functional.one_hot(torch.randint(0, 10, (64,), device=dev), 10).float()
    m1 = mnist_cnn(seed=1).to(dev)
    e1 = StepEngine(m1, optimizer='adam', loss='categorical_crossentropy', device=dev, use_cuda_graph=False)
    for _ in range(5):
        e1.step(x, y)
    path = str(tmp_path / 'cnn')
    save_checkpoint(path, m1, e1.optimizer, step=5, rank=0)
    m2 = mnist_cnn(seed=2).to(dev)
    e2 = StepEngine(m2, optimizer='adam', loss='categorical_crossentropy', device=dev, use_cuda_graph=False)
    e2.step(x, y)                                       # builds the fused step and its shadows from the WRONG weights
    assert e2._direct is not None
    load_checkpoint(path, m2, e2.optimizer)
    l1, l2 = float(e1.step(x, y)), float(e2.step(x, y))
    assert abs(l1 - l2) < 1e-4 * max(1.0, abs(l1)), (l1, l2)
    for a, b in zip(m1.get_weights(), m2.get_weights()):
        np.testing.assert_allclose(a, b, rtol=0, atol=2e-5)
    assert int(e1.optimizer.step_count.item()) == int(e2.optimizer.step_count.item()) == 6


def test_uncovered_model_on_cuda_raises_instead_of_falling_back(monkeypatch):
    """'No silent PyTorch fallback on a GPU box': a model no hand-written kernel covers (GRU(64) over 3 input features,
    the LEN_EXTRA_IN > 1 case of /root/reference/ddl_nyiso_aztk.py:80-88) must raise, and train eagerly only when asked."""
    from distributeddeeplearningspark_b200.models import Sequential
    from distributeddeeplearningspark_b200.models.layers import GRU, Dense
    from distributeddeeplearningspark_b200.ops.fused import NativePathError
    from distributeddeeplearningspark_b200.trainers import StepEngine
    dev = torch.device('cuda', 0)

    def make():
        m = Sequential([GRU(64, input_shape=(25, 3)), Dense(1)])
        m.compile(optimizer='adagrad', loss='mean_squared_error')
        return m.to(dev)
    x, y = torch.rand(32, 25, 3, device=dev), torch.rand(32, 1, device=dev)
    monkeypatch.delenv('DDLS_ALLOW_EAGER', raising=False)
    with pytest.raises(NativePathError, match='DDLS_ALLOW_EAGER'):
        StepEngine(make(), optimizer='adagrad', loss='mean_squared_error', device=dev, use_cuda_graph=False).step(x, y)
    monkeypatch.setenv('DDLS_ALLOW_EAGER', '1')
    eng = StepEngine(make(), optimizer='adagrad', loss='mean_squared_error', device=dev, use_cuda_graph=False)
    l0 = float(eng.step(x, y))
    for _ in range(10):
        l1 = float(eng.step(x, y))
    assert l1 < l0
