"""Lag/lead sliding-window featuriser for the NYISO forecaster.

Reproduces /root/reference/ddl_nyiso_aztk.py:139-177: per ``Name`` partition
ordered by ``TimeStamp``, 24 lag columns of the normalised load (oldest first),
the next hour's normalised load (target), the next hour's load in original
units and the next hour's normalised temperature (exogenous input); rows with
nulls are dropped; features = 24 lags + next-hour temperature reshaped (25,1).
"""
from __future__ import annotations

from typing import Optional, Sequence

import numpy as np

from .frame import DataFrame, VectorAssembler
from .transformers import ReshapeTransformer


def _shift(col: np.ndarray, n: int, groups: Optional[np.ndarray]) -> np.ndarray:
    """n > 0: lag by n (value from n rows earlier); n < 0: lead.  NaN where undefined or across groups."""
    x = np.asarray(col, dtype=np.float64)
    out = np.full_like(x, np.nan)
    if n > 0:
        out[n:] = x[:-n]
        if groups is not None:
            bad = np.ones(len(x), dtype=bool)
            bad[n:] = groups[n:] != groups[:-n]
            out[bad] = np.nan
    elif n < 0:
        k = -n
        out[:-k] = x[k:]
        if groups is not None:
            bad = np.ones(len(x), dtype=bool)
            bad[:-k] = groups[:-k] != groups[k:]
            out[bad] = np.nan
    else:
        out = x.copy()
    return out


class Window:
    """``Window.partitionBy(col).orderBy(col)`` analogue."""

    def __init__(self, partition_by: Optional[str] = None, order_by: Optional[str] = None):
        self.partition_by, self.order_by = partition_by, order_by

    @staticmethod
    def partitionBy(col: str) -> 'Window':
        return Window(partition_by=col)

    def orderBy(self, col: str) -> 'Window':
        return Window(self.partition_by, col)

    def sorted(self, df: DataFrame) -> DataFrame:
        keys = []
        if self.order_by is not None:
            keys.append(df[self.order_by])
        if self.partition_by is not None:
            keys.append(df[self.partition_by])
        if not keys:
            return df
        idx = np.lexsort(tuple(keys))
        return df._take(idx)

    def groups(self, df: DataFrame):
        if self.partition_by is None:
            return None
        _, inv = np.unique(df[self.partition_by], return_inverse=True)
        return inv


def lag(df: DataFrame, col: str, count: int, window: Window, output_col: str) -> DataFrame:
    """``df.withColumn(out, F.lag(col, count).over(window))``; ``df`` must already be window-sorted."""
    return df.withColumn(output_col, _shift(df[col], count, window.groups(df)))


def lead(df: DataFrame, col: str, count: int, window: Window, output_col: str) -> DataFrame:
    return df.withColumn(output_col, _shift(df[col], -count, window.groups(df)))


def make_nyiso_windows(df: DataFrame, len_seq_in: int = 24, len_seq_out: int = 1, norm_load: str = 'NormLoad',
                       norm_temp: str = 'NormTemp', orig_load: str = 'HourAvgLoad', name_col: str = 'Name',
                       ts_col: str = 'TimeStamp') -> DataFrame:
    """Full featuriser (ddl_nyiso_aztk.py:139-177).  Returns a frame with ``features``,
    ``feature_matrix`` (len_seq_in+len_seq_out, 1), ``labels``, ``labels2``, ``label_matrix``."""
    w = Window.partitionBy(name_col).orderBy(ts_col)
    df = w.sorted(df)
    for n in range(len_seq_in, 0, -1):
        df = lag(df, norm_load, n, w, f'{norm_load}_lag{n}')
    for n in range(1, len_seq_out + 1):
        df = lead(df, norm_load, n, w, f'{norm_load}_next{n}')
        df = lead(df, orig_load, n, w, f'Orig{norm_load}_next{n}')
        df = lead(df, norm_temp, n, w, f'{norm_temp}_next{n}')
    df = df.na.drop()
    feats = [f'{norm_load}_lag{n}' for n in range(len_seq_in, 0, -1)] + \
            [f'{norm_temp}_next{n}' for n in range(1, len_seq_out + 1)]
    df = VectorAssembler(feats, 'features').transform(df)
    df = ReshapeTransformer('features', 'feature_matrix', (len_seq_in + len_seq_out, 1)).transform(df)
    df = VectorAssembler([f'{norm_load}_next{n}' for n in range(1, len_seq_out + 1)], 'labels').transform(df)
    df = VectorAssembler([f'Orig{norm_load}_next{n}' for n in range(1, len_seq_out + 1)], 'labels2').transform(df)
    df = ReshapeTransformer('labels', 'label_matrix', (len_seq_out, 1)).transform(df)
    return df


def split_train_test(df: DataFrame, len_test: int, ts_col: str = 'TimeStamp'):
    """ddl_nyiso_aztk.py:179-183: train = all but the last ``len_test`` rows, test = last rows ascending."""
    df = df.orderBy(ts_col)
    n = df.count()
    return df.limit(n - len_test), df.orderBy(ts_col, ascending=False).limit(len_test).orderBy(ts_col)
