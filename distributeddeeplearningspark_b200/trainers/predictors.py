"""``ModelPredictor``, ``AccuracyEvaluator`` and ``get_MAPE``.

Upstream ``ModelPredictor.predict`` is ``df.rdd.mapPartitions(_predict)``: every executor deserialises the
model once per partition and then calls ``model.predict`` **row by row** (batch = 1).  Here
the same fan-out runs through ``spark.run`` — one rank (GPU) per partition, large device batches — in two
forms:

* driver form (``num_workers=N``): the driver stages the feature partitions, launches an N-rank barrier stage,
  every rank predicts its partition and the predictions come back through the stage's result channel (the
  ``collect()`` of the reference) in row order;
* in-job form (called by a rank of a running job): every rank predicts ``partition_indices(rank, size)`` and
  the predictions are all-gathered, so each rank ends up with the full frame.

``AccuracyEvaluator`` / ``evaluate_accuracy`` reduce (correct, total) counters over the ranks instead of
gathering predictions.  API per /root/reference/ddl_mnist_aztk.py:202-210 and
/root/reference/ddl_nyiso_aztk.py:220-246.
"""
from __future__ import annotations

import os
import shutil
import tempfile
from typing import Optional

import numpy as np
import torch

from .. import runtime
from ..data.frame import DataFrame
from ..models.sequential import Sequential


class Predictor:
    def predict(self, dataframe: DataFrame) -> DataFrame:  # pragma: no cover - interface
        raise NotImplementedError


def _predict_batches(model: Sequential, x: np.ndarray, batch_size: int, dev: torch.device) -> np.ndarray:
    model = model.to(dev).eval()
    outs = []
    xt = torch.from_numpy(np.array(x, dtype=np.float32, order='C'))       # a writable copy (partitions may be mmap'ed)
    pin = dev.type == 'cuda'
    with torch.no_grad():
        for lo in range(0, xt.shape[0], batch_size):
            xb = xt[lo:lo + batch_size]
            if pin:
                xb = xb.pin_memory().to(dev, non_blocking=True)
            outs.append(model(xb).float().cpu())
    if not outs:
        return np.zeros((0,) + tuple(model.output_shape), dtype=np.float64)
    return torch.cat(outs).numpy().astype(np.float64)


def _partition_worker(spec: dict) -> dict:
    """Runs on every rank of a prediction stage: predict my partition; optionally count arg-max hits against the labels
    and reduce (correct, total) over the ranks (``AccuracyEvaluator`` 'in a distributed manner')."""
    from ..models.sequential import deserialize_model
    runtime.init()
    rank, size, dev = runtime.rank(), runtime.size(), runtime.device()
    model = deserialize_model(spec['model'])
    x = np.load(os.path.join(spec['parts'], f'x{rank}.npy'), mmap_mode='r')
    pred = _predict_batches(model, np.asarray(x), spec['batch_size'], dev)
    out = {'rank': rank, 'rows': int(pred.shape[0])}
    if spec.get('count_hits'):
        y = np.load(os.path.join(spec['parts'], f'y{rank}.npy')).reshape(-1).astype(np.int64)
        hits = float(np.sum(pred.reshape(pred.shape[0], -1).argmax(axis=1) == y))
        t = torch.tensor([hits, float(pred.shape[0])], dtype=torch.float32, device=dev)
        if size > 1:
            from ..parallel.comm import Sum
            runtime.communicator().allreduce_(t, Sum)
        out['correct'], out['total'] = float(t[0]), float(t[1])
    else:
        out['prediction'] = pred
    return out


class ModelPredictor(Predictor):
    def __init__(self, keras_model: Sequential, features_col: str = 'features', output_col: str = 'prediction',
                 batch_size: int = 4096, device=None, num_workers: Optional[int] = None, backend=None):
        self.model = keras_model
        self.features_column = features_col
        self.output_column = output_col
        self.batch_size = int(batch_size)
        self.device = device
        self.num_workers = num_workers
        self.backend = backend

    def _device(self) -> torch.device:
        if self.device is not None:
            return torch.device(self.device)
        if runtime.is_initialized():
            return runtime.device()
        return torch.device('cuda', torch.cuda.current_device()) if torch.cuda.is_available() else torch.device('cpu')

    def predict_array(self, x: np.ndarray) -> np.ndarray:
        return _predict_batches(self.model, x, self.batch_size, self._device())

    # -- driver form: one rank per partition through spark.run ----------------------------------------
    def _stage(self, dataframe: DataFrame, n: int, label_col: Optional[str] = None) -> str:
        base = '/dev/shm' if os.path.isdir('/dev/shm') else None
        d = tempfile.mkdtemp(prefix='ddls_pred_', dir=base)
        x = np.asarray(dataframe[self.features_column])
        y = None if label_col is None else np.asarray(dataframe[label_col])
        for i in range(n):
            idx = dataframe.partition_indices(i, n)
            np.save(os.path.join(d, f'x{i}.npy'), np.ascontiguousarray(x[idx], dtype=np.float32))
            if y is not None:
                np.save(os.path.join(d, f'y{i}.npy'), np.ascontiguousarray(y[idx]))
        return d

    def _run_stage(self, dataframe: DataFrame, label_col: Optional[str] = None):
        from .. import spark
        from ..models.sequential import serialize_model
        n = int(self.num_workers)
        parts = self._stage(dataframe, n, label_col)
        try:
            spec = {'model': serialize_model(self.model), 'parts': parts, 'batch_size': self.batch_size,
                    'count_hits': label_col is not None}
            return spark.run(_partition_worker, args=(spec,), num_proc=n, backend=self.backend)
        finally:
            shutil.rmtree(parts, ignore_errors=True)

    def predict(self, dataframe: DataFrame) -> DataFrame:
        n_rows = dataframe.count()
        if self.num_workers and int(self.num_workers) > 1 and not runtime.is_initialized():
            outs = self._run_stage(dataframe)
            n = int(self.num_workers)
            first = next(o['prediction'] for o in outs if o['rows'] > 0)
            pred = np.zeros((n_rows,) + first.shape[1:], dtype=np.float64)
            for o in outs:                                   # results arrive ordered by rank = partition id
                pred[dataframe.partition_indices(o['rank'], n)] = o['prediction']
        elif runtime.is_initialized() and runtime.size() > 1:
            rank, size = runtime.rank(), runtime.size()
            idx = dataframe.partition_indices(rank, size)
            x = np.asarray(dataframe[self.features_column])[idx]
            mine = self.predict_array(x)
            comm = runtime.communicator()
            lens = comm.allgather_object(int(mine.shape[0]))
            width = int(np.prod(mine.shape[1:])) if mine.ndim > 1 else 1
            pad = torch.zeros((max(lens), width), dtype=torch.float32, device=runtime.device())
            pad[:mine.shape[0]] = torch.from_numpy(mine.reshape(mine.shape[0], -1).astype(np.float32))
            allp = comm.allgather(pad).cpu().numpy().astype(np.float64)          # [size, max_len, width]
            pred = np.zeros((n_rows, width), dtype=np.float64)
            for r in range(size):
                pred[dataframe.partition_indices(r, size)] = allp[r, :lens[r]]
        else:
            pred = self.predict_array(np.asarray(dataframe[self.features_column]))
        return dataframe.withColumn(self.output_column, pred.reshape(pred.shape[0], -1))

    def count_hits(self, dataframe: DataFrame, label_col: str):
        """(correct, total) of arg-max predictions against an integer label column, counted partition by partition on
        ``num_workers`` ranks and summed with one allreduce — no prediction leaves its rank."""
        outs = self._run_stage(dataframe, label_col)
        return outs[0]['correct'], outs[0]['total']


class Evaluator:
    def evaluate(self, dataframe: DataFrame):  # pragma: no cover - interface
        raise NotImplementedError


class AccuracyEvaluator(Evaluator):
    """count(prediction == label) / count() (two Spark actions upstream; one pass here)."""

    def __init__(self, prediction_col: str = 'prediction_index', label_col: str = 'label'):
        self.prediction_column, self.label_column = prediction_col, label_col

    def evaluate(self, dataframe: DataFrame) -> float:
        p = np.asarray(dataframe[self.prediction_column]).reshape(-1)
        y = np.asarray(dataframe[self.label_column]).reshape(-1)
        n = p.shape[0]
        correct = float(np.sum(p.astype(np.int64) == y.astype(np.int64)))
        total = float(n)
        if runtime.is_initialized() and runtime.size() > 1:  # data-parallel evaluation: reduce the counters
            from ..parallel import allreduce, Sum
            t = allreduce(torch.tensor([correct, total], dtype=torch.float32, device=runtime.device()), op=Sum)
            correct, total = float(t[0]), float(t[1])
        return correct / total if total else float('nan')


def get_MAPE(actual, pred) -> float:
    """Mean absolute percentage error, inf -> nan (/root/reference/ddl_nyiso_aztk.py:234-242)."""
    actual, pred = np.array(actual, dtype=np.float64), np.array(pred, dtype=np.float64)
    with np.errstate(divide='ignore', invalid='ignore'):
        mape = np.mean(np.abs((actual - pred) / actual)) * 100
    if mape == np.inf:
        mape = np.nan
    return float(mape)


def evaluate_accuracy(model: Sequential, test_set: DataFrame, features: str = 'matrix', label: str = 'label',
                      nb_classes: int = 10, num_workers: Optional[int] = None, backend=None) -> float:
    """The reference's ``evaluate_accuracy`` helper (/root/reference/ddl_mnist_aztk.py:202-210): predictor -> arg-max
    label index -> accuracy.  With ``num_workers`` > 1 the three stages run inside one ``spark.run`` job, a partition per
    rank, and only the two counters cross ranks."""
    from ..data.transformers import LabelIndexTransformer
    df = test_set.select(features, label)
    if num_workers and int(num_workers) > 1 and not runtime.is_initialized():
        correct, total = ModelPredictor(keras_model=model, features_col=features, num_workers=num_workers,
                                        backend=backend).count_hits(df, label)
        return correct / total if total else float('nan')
    df = ModelPredictor(keras_model=model, features_col=features).predict(df)
    df = LabelIndexTransformer(output_dim=nb_classes).transform(df)
    return AccuracyEvaluator(prediction_col='prediction_index', label_col=label).evaluate(df)
