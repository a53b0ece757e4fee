"""Column transformers with the dist-keras API the reference scripts call.

``OneHotTransformer`` (/root/reference/ddl_mnist_aztk.py:126),
``MinMaxTransformer`` (ddl_mnist_aztk.py:133-136, ddl_nyiso_aztk.py:123-135,225-229),
``ReshapeTransformer`` (ddl_mnist_aztk.py:142, ddl_nyiso_aztk.py:159,175),
``DenseTransformer`` (ddl_mnist_aztk.py:151), ``LabelIndexTransformer``
(ddl_mnist_aztk.py:205).  Upstream each one is ``df.rdd.map(row_fn).toDF()``;
here each is a single vectorised pass over the column (and the MNIST
normalise/one-hot pair also exists as an on-device kernel, ``ops.preprocess``).
"""
from __future__ import annotations

from typing import Sequence

import numpy as np

from .frame import DataFrame


class Transformer:
    def transform(self, dataframe: DataFrame) -> DataFrame:  # pragma: no cover - interface
        raise NotImplementedError


class MinMaxTransformer(Transformer):
    """x -> (x - o_min) * (n_max - n_min) / (o_max - o_min) + n_min, scalar or vector column."""

    def __init__(self, n_min, n_max, o_min, o_max, input_col: str, output_col: str, is_vector: bool = True):
        self.n_min, self.n_max = float(n_min), float(n_max)
        self.o_min, self.o_max = float(o_min), float(o_max)
        self.input_col, self.output_col, self.is_vector = input_col, output_col, is_vector
        self.scale = (self.n_max - self.n_min) / (self.o_max - self.o_min)

    def apply(self, x: np.ndarray) -> np.ndarray:
        return (np.asarray(x, dtype=np.float64) - self.o_min) * self.scale + self.n_min

    def transform(self, dataframe):
        return dataframe.withColumn(self.output_col, self.apply(dataframe[self.input_col]))


class StandardTransformer(Transformer):
    """(x - mean) / std per column (upstream ``StandardTransformer``; not used by the scripts)."""

    def __init__(self, input_col: str, output_col: str, mean=None, std=None):
        self.input_col, self.output_col, self.mean, self.std = input_col, output_col, mean, std

    def transform(self, dataframe):
        x = np.asarray(dataframe[self.input_col], dtype=np.float64)
        mean = x.mean(axis=0) if self.mean is None else self.mean
        std = x.std(axis=0) if self.std is None else self.std
        std = np.where(np.asarray(std) == 0, 1.0, std)
        return dataframe.withColumn(self.output_col, (x - mean) / std)


class OneHotTransformer(Transformer):
    def __init__(self, output_dim: int, input_col: str, output_col: str):
        self.output_dim, self.input_col, self.output_col = int(output_dim), input_col, output_col

    def transform(self, dataframe):
        idx = np.asarray(dataframe[self.input_col]).astype(np.int64)
        if idx.size and (idx.min() < 0 or idx.max() >= self.output_dim):
            raise ValueError('label outside [0, output_dim)')
        out = np.zeros((idx.shape[0], self.output_dim), dtype=np.float64)
        out[np.arange(idx.shape[0]), idx] = 1.0
        return dataframe.withColumn(self.output_col, out)


class BinaryLabelTransformer(Transformer):
    def __init__(self, input_column: str, output_column: str, label):
        self.input_column, self.output_column, self.label = input_column, output_column, label

    def transform(self, dataframe):
        return dataframe.withColumn(self.output_column,
                                    (np.asarray(dataframe[self.input_column]) == self.label).astype(np.float64))


class ReshapeTransformer(Transformer):
    def __init__(self, input_col: str, output_col: str, shape: Sequence[int]):
        self.input_col, self.output_col, self.shape = input_col, output_col, tuple(int(s) for s in shape)

    def transform(self, dataframe):
        x = np.asarray(dataframe[self.input_col])
        return dataframe.withColumn(self.output_col, x.reshape((x.shape[0],) + self.shape))


class DenseTransformer(Transformer):
    """SparseVector -> DenseVector upstream; columns here are always dense, so this is a typed copy."""

    def __init__(self, input_col: str, output_col: str):
        self.input_col, self.output_col = input_col, output_col

    def transform(self, dataframe):
        return dataframe.withColumn(self.output_col, np.ascontiguousarray(dataframe[self.input_col], dtype=np.float64))


class LabelIndexTransformer(Transformer):
    """arg-max of the prediction vector -> ``prediction_index``."""

    def __init__(self, output_dim: int, input_col: str = 'prediction', output_col: str = 'prediction_index',
                 activation_threshold: float = 0.55):
        self.output_dim, self.input_col, self.output_col = int(output_dim), input_col, output_col
        self.activation_threshold = activation_threshold

    def transform(self, dataframe):
        p = np.asarray(dataframe[self.input_col])
        if p.ndim == 1 or p.shape[-1] == 1:  # single sigmoid output
            idx = (p.reshape(p.shape[0]) >= self.activation_threshold).astype(np.float64)
        else:
            idx = p.reshape(p.shape[0], -1).argmax(axis=1).astype(np.float64)
        return dataframe.withColumn(self.output_col, idx)
