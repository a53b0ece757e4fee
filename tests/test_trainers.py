import numpy as np
import pytest
import torch

from distributeddeeplearningspark_b200.data import (DataFrame, MinMaxTransformer, OneHotTransformer,
                                                     ReshapeTransformer, make_nyiso_windows, mnist_like_frame,
                                                     nyiso_like_frame, split_train_test)
from distributeddeeplearningspark_b200.models import mnist_cnn, nyiso_gru, nyiso_lstm
from distributeddeeplearningspark_b200.spark import LocalBarrierBackend
from distributeddeeplearningspark_b200.trainers import (ADAG, AccuracyEvaluator, AllreduceTrainer, DynSGD,
                                                         ModelPredictor, SingleTrainer, evaluate_accuracy, get_MAPE)

ENV_BACKEND = dict(timeout=600)


def _nyiso_frames(n_hours=700):
    df = nyiso_like_frame(n_hours, seed=1)
    lo, hi = df.agg_min('HourAvgLoad'), df.agg_max('HourAvgLoad')
    df = MinMaxTransformer(0, 1, df.agg_min('temperature'), df.agg_max('temperature'), 'temperature', 'NormTemp',
                           is_vector=False).transform(df)
    df = MinMaxTransformer(0, 1, lo, hi, 'HourAvgLoad', 'NormLoad', is_vector=False).transform(df)
    tr, te = split_train_test(make_nyiso_windows(df), 120)
    return tr, te, lo, hi


def test_single_trainer_learns_and_predicts_mape(monkeypatch):
    monkeypatch.setenv('DDLS_FORCE_CPU', '1')
    tr, te, lo, hi = _nyiso_frames()
    model = nyiso_gru(units=16, seed=0)
    t = SingleTrainer(keras_model=model, worker_optimizer='adam', loss='mean_squared_error', num_epoch=6,
                      batch_size=32, features_col='feature_matrix', label_col='labels',
                      optimizer_kwargs={'lr': 0.01})
    trained = t.train(tr)
    h = t.get_history()[0]
    assert len(h) == (tr.count() * 6) // 32
    assert np.mean(h[-10:]) < 0.5 * np.mean(h[:10])
    assert t.get_training_time() > 0 and t.parameter_server.num_updates == len(h)
    pred = ModelPredictor(keras_model=trained, features_col='feature_matrix').predict(te.limit(24))
    pred = MinMaxTransformer(n_min=lo, n_max=hi, o_min=0.0, o_max=1.0, input_col='prediction',
                             output_col='prediction2', is_vector=True).transform(pred)
    mape = get_MAPE(pred['labels2'].tolist(), pred['prediction2'].tolist())
    assert 0 <= mape < 25.0


def _update_count(rows_per_worker, epochs, batch, window):
    return sum(((r * epochs) // batch) // window for r in rows_per_worker)


def test_update_count_formula_reproduces_notebook_1425():
    # 4 workers x 20 epochs x window 5 x batch 32 -> 1425 (ddl_nyiso_hdi.ipynb:608,817)
    assert _update_count([2856, 2850, 2850, 2850], 20, 32, 5) == 1425


@pytest.mark.parametrize('cls', [ADAG, DynSGD])
def test_adag_two_workers_gloo_update_counter_and_center(cls):
    tr, te, lo, hi = _nyiso_frames(560)                     # 415 training windows
    n = tr.count()
    model = nyiso_lstm(units=8, seed=0)
    t = cls(keras_model=model, worker_optimizer='adam', loss='mean_squared_error', num_workers=2, batch_size=16,
            communication_window=5, num_epoch=2, features_col='feature_matrix', label_col='labels',
            backend=LocalBarrierBackend(**ENV_BACKEND), seed=0)
    t.backend_env = None
    import os
    os.environ['DDLS_FORCE_CPU'] = '1'
    try:
        trained = t.train(tr)
    finally:
        os.environ.pop('DDLS_FORCE_CPU', None)
    rows = [len(tr.partition_indices(i, 2)) for i in range(2)]
    assert t.parameter_server.num_updates == _update_count(rows, 2, 16, 5)
    assert len(t.get_history()) == 2 and all(len(h) == (r * 2) // 16 for h, r in zip(t.get_history(), rows))
    w0 = model.get_weights()
    w1 = trained.get_weights()
    assert any(np.abs(a - b).max() > 1e-4 for a, b in zip(w0, w1))     # center moved
    assert all(np.isfinite(a).all() for a in w1)


@pytest.mark.parametrize('name', ['DOWNPOUR', 'AEASGD', 'EAMSGD', 'AveragingTrainer', 'EnsembleTrainer'])
def test_other_distkeras_trainers_two_workers_gloo(name):
    """The rest of dist-keras' trainer family on two gloo ranks: the model moves, stays finite, and the bookkeeping
    (update counter / returned models) has the documented shape."""
    import os
    from distributeddeeplearningspark_b200 import trainers as T
    tr, te, lo, hi = _nyiso_frames(400)
    model = nyiso_lstm(units=8, seed=0)
    kw = dict(keras_model=model, worker_optimizer='adagrad', loss='mean_squared_error', batch_size=16, num_epoch=2,
              features_col='feature_matrix', label_col='labels', backend=LocalBarrierBackend(**ENV_BACKEND), seed=0)
    if name == 'EnsembleTrainer':
        t = T.EnsembleTrainer(num_models=2, **kw)
    elif name == 'AveragingTrainer':
        t = T.AveragingTrainer(num_workers=2, **kw)
    else:
        t = getattr(T, name)(num_workers=2, communication_window=4, **kw)
    os.environ['DDLS_FORCE_CPU'] = '1'
    try:
        out = t.train(tr)
    finally:
        os.environ.pop('DDLS_FORCE_CPU', None)
    w0 = model.get_weights()
    models = out if name == 'EnsembleTrainer' else [out]
    assert len(models) == (2 if name == 'EnsembleTrainer' else 1)
    for m in models:
        w1 = m.get_weights()
        assert any(np.abs(a - b).max() > 1e-4 for a, b in zip(w0, w1)) and all(np.isfinite(a).all() for a in w1)
    if name == 'EnsembleTrainer':
        assert any(np.abs(a - b).max() > 1e-6 for a, b in zip(models[0].get_weights(), models[1].get_weights()))
    rows = [len(tr.partition_indices(i, 2)) for i in range(2)]
    if name in ('DOWNPOUR', 'AEASGD', 'EAMSGD'):
        assert t.parameter_server.num_updates == _update_count(rows, 2, 16, 4)
    if name == 'AveragingTrainer':
        assert t.parameter_server.num_updates == 2
    assert len(t.get_history()) == 2 and all(len(h) > 0 for h in t.get_history())


def test_allreduce_trainer_mnist_two_ranks_and_accuracy_eval():
    df = mnist_like_frame(640, seed=0)
    df = OneHotTransformer(10, 'label', 'label_encoded').transform(df)
    df = MinMaxTransformer(0.0, 1.0, 0.0, 250.0, 'features', 'features_normalized').transform(df)
    df = ReshapeTransformer('features_normalized', 'matrix', (28, 28, 1)).transform(df)
    import os
    os.environ['DDLS_FORCE_CPU'] = '1'
    try:
        t = AllreduceTrainer(keras_model=mnist_cnn(seed=0), worker_optimizer='adam',
                             loss='categorical_crossentropy', num_workers=2, batch_size=32, num_epoch=2,
                             features_col='matrix', label_col='label_encoded',
                             backend=LocalBarrierBackend(**ENV_BACKEND))
        trained = t.train(df)
        acc = evaluate_accuracy(trained, df.limit(200), features='matrix')
    finally:
        os.environ.pop('DDLS_FORCE_CPU', None)
    assert acc > 0.5                                         # synthetic classes are separable
    h = t.get_averaged_history()
    assert h[-1] < h[0]


def test_accuracy_evaluator_counts():
    df = DataFrame({'prediction_index': np.array([1.0, 2, 3, 4]), 'label': np.array([1, 2, 0, 4])})
    assert AccuracyEvaluator('prediction_index', 'label').evaluate(df) == 0.75


def test_step_engine_run_stream_cpu_path():
    """run_stream on CPU degrades to a plain loop over the host batches (same losses as step())."""
    import torch
    import distributeddeeplearningspark_b200 as ddls
    from distributeddeeplearningspark_b200.trainers import StepEngine
    torch.manual_seed(0)
    xs = [torch.rand(8, 25, 1) for _ in range(4)]
    ys = [torch.rand(8, 1) for _ in range(4)]
    ma, mb = ddls.models.nyiso_gru(seed=2), ddls.models.nyiso_gru(seed=2)
    ea = StepEngine(ma, optimizer='adagrad', loss='mean_squared_error', device='cpu')
    eb = StepEngine(mb, optimizer='adagrad', loss='mean_squared_error', device='cpu')
    la = ea.run_stream(zip(xs, ys), xs[0], ys[0])
    lb = [float(eb.step(x, y)) for x, y in zip(xs, ys)]
    assert la == lb and la[-1] < la[0] * 1.5


def _cnn_frame(n):
    df = mnist_like_frame(n, seed=3)
    df = MinMaxTransformer(0.0, 1.0, 0.0, 250.0, 'features', 'features_normalized').transform(df)
    return ReshapeTransformer('features_normalized', 'matrix', (28, 28, 1)).transform(df)


def test_model_predictor_fans_out_over_partitions_like_map_partitions(monkeypatch):
    """ModelPredictor(num_workers=3): a partition per rank through spark.run, predictions back in row order
    (/root/reference/ddl_mnist_aztk.py:204-207), and evaluate_accuracy reducing only (correct, total) counters."""
    monkeypatch.setenv('DDLS_FORCE_CPU', '1')
    df = _cnn_frame(101)                                     # partitions of 34 / 34 / 33 rows
    model = mnist_cnn(seed=5)
    one = ModelPredictor(keras_model=model, features_col='matrix').predict(df)
    many = ModelPredictor(keras_model=model, features_col='matrix', num_workers=3, batch_size=16,
                          backend=LocalBarrierBackend(**ENV_BACKEND)).predict(df)
    np.testing.assert_allclose(np.asarray(many['prediction']), np.asarray(one['prediction']), rtol=1e-5, atol=1e-6)
    a1 = evaluate_accuracy(model, df, features='matrix')
    a3 = evaluate_accuracy(model, df, features='matrix', num_workers=3, backend=LocalBarrierBackend(**ENV_BACKEND))
    assert abs(a1 - a3) < 1e-9


def _predict_inside_job(n_rows):
    import distributeddeeplearningspark_b200 as ddls
    ddls.init()
    df = _cnn_frame(n_rows)
    out = ModelPredictor(keras_model=mnist_cnn(seed=5), features_col='matrix', batch_size=8).predict(df)
    return np.asarray(out['prediction'])


def test_model_predictor_inside_a_running_job_allgathers_predictions(monkeypatch):
    from distributeddeeplearningspark_b200 import spark
    monkeypatch.setenv('DDLS_FORCE_CPU', '1')
    outs = spark.run(_predict_inside_job, args=(37,), num_proc=2, backend=LocalBarrierBackend(**ENV_BACKEND),
                     env={'DDLS_FORCE_CPU': '1'})
    ref = np.asarray(ModelPredictor(keras_model=mnist_cnn(seed=5), features_col='matrix').predict(_cnn_frame(37))['prediction'])
    for o in outs:                                           # every rank holds the whole frame's predictions
        np.testing.assert_allclose(o, ref, rtol=1e-5, atol=1e-6)


def test_adag_trainer_reports_the_notebooks_1425_updates():
    """The notebook's recorded invariant, driven through ADAG.train(): 4 workers x 20 epochs x batch 32 x window 5 report
    1425 parameter-server updates (/root/reference/ddl_nyiso_hdi.ipynb:608, 817).  The count is
    sum_i floor(floor(rows_i * 20 / 32) / 5); Spark's exact partition sizes are not recorded, so the hourly series is sized
    (11,566 hours -> 11,421 training windows after 24 lags, 1 lead and the 120-row test split) such that this framework's
    balanced partitions (2856 + 3 x 2855 rows) land on the recorded number.  Tiny recurrent width: only the bookkeeping
    is under test."""
    import os
    tr, te, lo, hi = _nyiso_frames(11566)
    assert tr.count() == 11421 and te.count() == 120
    rows = [len(tr.partition_indices(i, 4)) for i in range(4)]
    assert rows == [2856, 2855, 2855, 2855] and _update_count(rows, 20, 32, 5) == 1425
    t = ADAG(keras_model=nyiso_lstm(units=2, seed=0), worker_optimizer='adam', loss='mean_squared_error', num_workers=4,
             batch_size=32, communication_window=5, num_epoch=20, features_col='feature_matrix', label_col='labels',
             backend=LocalBarrierBackend(**ENV_BACKEND), seed=0)
    os.environ['DDLS_FORCE_CPU'] = '1'
    try:
        t.train(tr)
    finally:
        os.environ.pop('DDLS_FORCE_CPU', None)
    assert t.parameter_server.num_updates == 1425
    assert [len(h) for h in t.get_history()] == [(r * 20) // 32 for r in rows]
    assert t.get_training_time() > 0


def test_trainer_observability_phase_times_metrics_jsonl_and_timeline(tmp_path, monkeypatch):
    """DDLS_METRICS / DDLS_TIMELINE sinks and the phase timers are wired into the worker loop (they existed
    without a call site)."""
    import json
    monkeypatch.setenv('DDLS_FORCE_CPU', '1')
    monkeypatch.setenv('DDLS_METRICS', str(tmp_path / 'metrics.jsonl'))
    monkeypatch.setenv('DDLS_METRICS_EVERY', '5')
    monkeypatch.setenv('DDLS_TIMELINE', str(tmp_path / 'timeline.json'))
    tr, _, _, _ = _nyiso_frames(300)
    t = SingleTrainer(keras_model=nyiso_gru(units=4, seed=0), worker_optimizer='adagrad', loss='mean_squared_error',
                      num_epoch=2, batch_size=16, features_col='feature_matrix', label_col='labels')
    t.train(tr)
    steps = len(t.get_history()[0])
    ph = t.get_phase_times()[0]
    assert ph['step']['count'] == steps and ph['feed']['count'] >= steps and ph['step']['total_s'] > 0
    rows = [json.loads(l) for l in open(tmp_path / 'metrics.jsonl')]
    assert [r['step'] for r in rows[:-1]] == list(range(5, steps + 1, 5)) and rows[-1].get('final') and rows[-1]['step'] == steps
    assert all(r['samples_per_s'] > 0 for r in rows)
    ev = json.load(open(tmp_path / 'timeline.json'))['traceEvents']
    assert sum(1 for e in ev if e['name'] == 'step') == steps and any(e['ph'] == 'i' for e in ev)


def test_trainer_streams_arrow_partitions_with_the_same_result_as_arrays(monkeypatch):
    """DDLS_PARTITION_FORMAT=arrow: partitions are staged as Arrow IPC record batches and streamed through the pinned ring
    (the mapInArrow-shaped feed); the training run is the same as with memory-mapped arrays."""
    monkeypatch.setenv('DDLS_FORCE_CPU', '1')
    tr, _, _, _ = _nyiso_frames(420)
    hist = {}
    for fmt in ('npy', 'arrow'):
        monkeypatch.setenv('DDLS_PARTITION_FORMAT', fmt)
        monkeypatch.setenv('DDLS_ARROW_BATCH_ROWS', '50')
        t = SingleTrainer(keras_model=nyiso_lstm(units=4, seed=1), worker_optimizer='adam', loss='mean_squared_error',
                          num_epoch=2, batch_size=16, features_col='feature_matrix', label_col='labels', seed=0)
        t.train(tr)
        hist[fmt] = (t.get_history()[0], t.parameter_server.num_updates)
    assert hist['npy'][1] == hist['arrow'][1] and len(hist['npy'][0]) == (tr.count() * 2) // 16
    np.testing.assert_allclose(hist['arrow'][0], hist['npy'][0], rtol=1e-6, atol=1e-7)
