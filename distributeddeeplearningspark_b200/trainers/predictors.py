"""``ModelPredictor``, ``AccuracyEvaluator`` and ``get_MAPE``.

Upstream ``ModelPredictor.predict`` deserialises the model once per partition and then calls
``model.predict`` **row by row** (batch = 1); here each partition is predicted in
large device batches (and, when a job is running, each rank takes its own partitions and the
(correct, total) counters are allreduced).  API per /root/reference/ddl_mnist_aztk.py:202-210 and
/root/reference/ddl_nyiso_aztk.py:220-246.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from .. import runtime
from ..data.frame import DataFrame
from ..models.sequential import Sequential


class Predictor:
    def predict(self, dataframe: DataFrame) -> DataFrame:  # pragma: no cover - interface
        raise NotImplementedError


class ModelPredictor(Predictor):
    def __init__(self, keras_model: Sequential, features_col: str = 'features', output_col: str = 'prediction',
                 batch_size: int = 4096, device=None):
        self.model = keras_model
        self.features_column = features_col
        self.output_column = output_col
        self.batch_size = int(batch_size)
        self.device = device

    def _device(self) -> torch.device:
        if self.device is not None:
            return torch.device(self.device)
        if runtime.is_initialized():
            return runtime.device()
        return torch.device('cuda', torch.cuda.current_device()) if torch.cuda.is_available() else torch.device('cpu')

    @torch.no_grad()
    def predict_array(self, x: np.ndarray) -> np.ndarray:
        dev = self._device()
        model = self.model.to(dev).eval()
        outs = []
        xt = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
        pin = dev.type == 'cuda'
        for lo in range(0, xt.shape[0], self.batch_size):
            xb = xt[lo:lo + self.batch_size]
            if pin:
                xb = xb.pin_memory().to(dev, non_blocking=True)
            outs.append(model(xb).float().cpu())
        if not outs:
            return np.zeros((0,) + tuple(self.model.output_shape), dtype=np.float64)
        return torch.cat(outs).numpy().astype(np.float64)

    def predict(self, dataframe: DataFrame) -> DataFrame:
        x = np.asarray(dataframe[self.features_column])
        pred = self.predict_array(x)
        return dataframe.withColumn(self.output_column, pred.reshape(pred.shape[0], -1))


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
                      nb_classes: int = 10) -> float:
    """The reference's ``evaluate_accuracy`` helper (/root/reference/ddl_mnist_aztk.py:202-210)."""
    from ..data.transformers import LabelIndexTransformer
    df = test_set.select(features, label)
    df = ModelPredictor(keras_model=model, features_col=features).predict(df)
    df = LabelIndexTransformer(output_dim=nb_classes).transform(df)
    return AccuracyEvaluator(prediction_col='prediction_index', label_col=label).evaluate(df)
