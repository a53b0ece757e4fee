import numpy as np
import pytest
import torch

from distributeddeeplearningspark_b200.data import (DataFrame, DenseTransformer, DeviceFeeder, LabelIndexTransformer,
                                                     MinMaxTransformer, OneHotTransformer, ReshapeTransformer,
                                                     VectorAssembler, make_nyiso_windows, mnist_like_frame,
                                                     nyiso_like_frame, split_train_test)
from distributeddeeplearningspark_b200.trainers import get_MAPE

# first three rows of the NYISO notebook and their normalised values (ddl_nyiso_hdi.ipynb:207-209,237-239)
RAW = [(4836.967, 34.44), (4613.692, 33.79), (4448.508, 33.13)]
NORM = [(0.13131163940862922, 0.3608378292605522), (0.09975432572999143, 0.3539617052787475),
        (0.07640749647148892, 0.34697979477414576)]


def _range_from_pairs(x0, n0, x1, n1):
    slope = (n1 - n0) / (x1 - x0)
    o_min = x0 - n0 / slope
    return o_min, o_min + 1.0 / slope


def test_minmax_reproduces_notebook_golden_rows():
    lo, hi = _range_from_pairs(RAW[0][0], NORM[0][0], RAW[1][0], NORM[1][0])
    tlo, thi = _range_from_pairs(RAW[0][1], NORM[0][1], RAW[1][1], NORM[1][1])
    df = DataFrame({'HourAvgLoad': [r[0] for r in RAW], 'temperature': [r[1] for r in RAW]})
    df = MinMaxTransformer(0.0, 1.0, tlo, thi, 'temperature', 'NormTemp', is_vector=False).transform(df)
    df = MinMaxTransformer(0.0, 1.0, lo, hi, 'HourAvgLoad', 'NormLoad', is_vector=False).transform(df)
    assert df['NormLoad'][2] == pytest.approx(NORM[2][0], abs=1e-9)     # third row is the held-out check
    assert df['NormTemp'][2] == pytest.approx(NORM[2][1], abs=1e-9)
    # inverse transform (ddl_nyiso_aztk.py:225-230) round-trips to original units
    inv = MinMaxTransformer(lo, hi, 0.0, 1.0, 'NormLoad', 'back', is_vector=True).transform(df)
    np.testing.assert_allclose(inv['back'], [r[0] for r in RAW], rtol=1e-9)


def test_mnist_minmax_uses_250_like_the_reference():
    df = DataFrame({'features': np.array([[0.0, 250.0, 255.0]])})
    out = MinMaxTransformer(0.0, 1.0, 0.0, 250.0, 'features', 'n').transform(df)
    np.testing.assert_allclose(out['n'][0], [0.0, 1.0, 1.02])     # ddl_mnist_aztk.py:133-134


def test_onehot_reshape_dense_labelindex():
    df = DataFrame({'label': np.array([3, 0, 9]), 'features': np.arange(3 * 784, dtype=np.float64).reshape(3, 784)})
    df = OneHotTransformer(10, 'label', 'label_encoded').transform(df)
    assert df['label_encoded'].shape == (3, 10) and df['label_encoded'][0, 3] == 1 and df['label_encoded'].sum() == 3
    df = ReshapeTransformer('features', 'matrix', (28, 28, 1)).transform(df)
    assert df['matrix'].shape == (3, 28, 28, 1) and df['matrix'][1, 0, 1, 0] == 785
    df = DenseTransformer('features', 'dense').transform(df)
    assert df['dense'].dtype == np.float64
    df = df.withColumn('prediction', np.eye(10)[[3, 1, 9]])
    df = LabelIndexTransformer(10).transform(df)
    np.testing.assert_array_equal(df['prediction_index'], [3, 1, 9])


def test_window_featuriser_layout_matches_notebook_row():
    """First surviving row: lag24 = first raw row's NormLoad, oldest lag first, target = lead by one
    (ddl_nyiso_hdi.ipynb:298)."""
    n = 60
    load = np.arange(n, dtype=np.float64) * 10 + 1000
    temp = np.arange(n, dtype=np.float64) * 0.5 + 30
    ts = np.datetime64('2016-01-02T00:00:00') + np.arange(n) * np.timedelta64(3600, 's')
    df = DataFrame({'TimeStamp': ts[::-1].copy(), 'Name': np.array(['N.Y.C.'] * n), 'HourAvgLoad': load[::-1].copy(),
                    'temperature': temp[::-1].copy()})               # deliberately unsorted
    df = MinMaxTransformer(0, 1, temp.min(), temp.max(), 'temperature', 'NormTemp', is_vector=False).transform(df)
    df = MinMaxTransformer(0, 1, load.min(), load.max(), 'HourAvgLoad', 'NormLoad', is_vector=False).transform(df)
    w = make_nyiso_windows(df)
    assert w.count() == n - 24 - 1
    first = w.limit(1)
    assert first['TimeStamp'][0] == np.datetime64('2016-01-03T00:00:00')
    nl = (load - load.min()) / (load.max() - load.min())
    nt = (temp - temp.min()) / (temp.max() - temp.min())
    np.testing.assert_allclose(first['features'][0][:24], nl[0:24])          # lag24 .. lag1 = oldest first
    assert first['features'][0][24] == pytest.approx(nt[25])                 # next hour's temperature
    assert first['labels'][0][0] == pytest.approx(nl[25])
    assert first['labels2'][0][0] == pytest.approx(load[25])
    assert first['feature_matrix'].shape == (1, 25, 1) and first['label_matrix'].shape == (1, 1, 1)
    assert w.schema['features'] == 'vector' and w.schema['feature_matrix'] == 'array<array<double>>'


def test_split_train_test_last_120_ascending():
    df = make_nyiso_windows(_norm(nyiso_like_frame(600)))
    tr, te = split_train_test(df, 120)
    assert te.count() == 120 and tr.count() == df.count() - 120
    assert np.all(np.diff(te['TimeStamp'].astype('int64')) > 0)
    assert te['TimeStamp'][0] > tr['TimeStamp'][-1]


def _norm(df):
    df = MinMaxTransformer(0, 1, df.agg_min('temperature'), df.agg_max('temperature'), 'temperature', 'NormTemp',
                           is_vector=False).transform(df)
    return MinMaxTransformer(0, 1, df.agg_min('HourAvgLoad'), df.agg_max('HourAvgLoad'), 'HourAvgLoad', 'NormLoad',
                             is_vector=False).transform(df)


def test_frame_ops():
    df = mnist_like_frame(50, as_columns=True)
    assert len(df.columns) == 785 and df.count() == 50
    feats = [c for c in df.columns if c != 'label']
    df = VectorAssembler(feats, 'features').transform(df).select('features', 'label')
    assert df['features'].shape == (50, 784)
    parts = df.repartition(8).partitions()
    assert sum(p.count() for p in parts) == 50 and max(p.count() for p in parts) - min(p.count() for p in parts) <= 1
    assert df.limit(7).count() == 7
    assert df.orderBy('label', ascending=False)['label'][0] == df['label'].max()
    rows = df.limit(2).collect()
    assert len(rows) == 2 and len(rows[0]['features']) == 784 and rows[0].label == rows[0]['label']
    lines = []
    df.show(3, print_fn=lines.append)
    assert 'only showing top 3 rows' in lines[-1]
    nan = DataFrame({'a': np.array([1.0, np.nan, 3.0]), 'b': np.array([[1.0, 2], [3, 4], [np.nan, 6]])})
    assert nan.na.drop().count() == 1


def test_get_mape():
    assert get_MAPE([[100.0], [200.0]], [[110.0], [180.0]]) == pytest.approx(10.0)
    assert np.isnan(get_MAPE([0.0, 1.0], [1.0, 1.0]))


def test_feeder_carry_over_counts_and_content():
    x = np.arange(10, dtype=np.float32).reshape(10, 1)
    f = DeviceFeeder({'x': x}, batch_size=4, device='cpu', num_epoch=2, carry_over=True)
    got = [b['x'].reshape(-1).tolist() for b in f]
    assert len(got) == len(f) == 5                       # 20 rows / 4
    assert got[2] == [8.0, 9.0, 0.0, 1.0]                # batch straddles the epoch boundary
    f2 = DeviceFeeder({'x': x}, batch_size=4, device='cpu', num_epoch=2)
    assert len(list(f2)) == 4                            # per-epoch drop_last


def test_distkeras_utils_surface():
    """The ``distkeras.utils`` helpers of SURVEY.md §2.2 E11."""
    import numpy as np
    import distributeddeeplearningspark_b200 as ddls
    from distributeddeeplearningspark_b200 import utils as U
    from distributeddeeplearningspark_b200.data.frame import DataFrame, Row
    r = U.new_dataframe_row(Row(a=1, b=2.0), 'c', [1, 2])
    assert r.a == 1 and r.c == [1, 2] and r.asDict() == {'a': 1, 'b': 2.0, 'c': [1, 2]}
    np.testing.assert_array_equal(U.to_one_hot_encoded_dense(2, 4), [0, 0, 1, 0])
    df = DataFrame({'x': np.arange(10.0), 'y': np.arange(10) % 2})
    assert U.precache(df, 2) == 10
    sh = U.shuffle(df, seed=3)
    assert sorted(np.asarray(sh['x']).tolist()) == list(range(10))
    m = ddls.models.nyiso_gru(seed=1)
    blob = U.serialize_keras_model(m)
    m2 = U.deserialize_keras_model(blob)
    for a, b in zip(m.get_weights(), m2.get_weights()):
        np.testing.assert_array_equal(a, b)
    assert U.history_executors_average([[1.0, 3.0, 5.0], [3.0, 5.0]]) == [2.0, 4.0]
    assert U.history_executor([[1.0], [2.0]], 1) == [2.0]
    t = ddls.trainers.ADAG(m, 'adagrad', 'mean_squared_error', num_workers=2)
    assert t.master_port == 5000 and t.parallelism_factor == 1 and t.communication_window == 12


def test_arrow_stream_feeder_cuts_the_same_minibatches_as_the_array_feeder(tmp_path):
    """Arrow record batches -> pinned ring: minibatches straddle record batches and epochs exactly like the array feeder's
    dist-keras-style carry-over batching; uint8 pixels stay uint8; the final partial minibatch is dropped."""
    import torch
    from distributeddeeplearningspark_b200.data import ArrowStreamFeeder, DeviceFeeder, ipc_file_source, write_partition_file
    rng = np.random.default_rng(0)
    x = rng.integers(0, 255, size=(103, 28, 28, 1), dtype=np.uint8)
    y = rng.random((103, 10)).astype(np.float32)
    path = str(tmp_path / 'part0.arrow')
    assert write_partition_file(path, {'x': x, 'y': y}, rows_per_batch=17) == 7
    source, rows, cols = ipc_file_source(path)
    assert rows == 103 and cols['x'] == ((28, 28, 1), np.dtype('uint8')) and cols['y'][0] == (10,)
    fa = ArrowStreamFeeder(source, {k: (k, s, d) for k, (s, d) in cols.items()}, 16, 'cpu', num_epoch=3, total_rows=rows)
    fb = DeviceFeeder({'x': x, 'y': y}, 16, 'cpu', num_epoch=3, carry_over=True)
    a, b = list(fa), [{k: v.clone() for k, v in d.items()} for d in fb]
    assert len(a) == len(b) == len(fa) == (103 * 3) // 16
    for da, db in zip(a, b):
        assert da['x'].dtype == torch.uint8 and torch.equal(da['x'], db['x']) and torch.equal(da['y'], db['y'])
