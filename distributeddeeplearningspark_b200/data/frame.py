"""A small columnar DataFrame with the Spark surface the reference scripts use.

The reference drives a Spark ``DataFrame`` (``select/repartition/cache/limit/
orderBy/na.drop/show/count/collect``, /root/reference/ddl_mnist_aztk.py:100-161,
/root/reference/ddl_nyiso_aztk.py:110-198) and every dist-keras transformer is
a per-row Python lambda over ``df.rdd``.  PySpark and a JVM are
not on this image, and per-row Python would starve a B200, so this class keeps
each column as one contiguous NumPy array (vector columns are 2-D+ arrays) and
all transforms are whole-column operations.  ``from_spark``/``from_arrow``
adapters accept real Spark or Arrow data when those are present.
"""
from __future__ import annotations

from typing import Callable, Dict, Iterable, List, Optional, Sequence, Union

import numpy as np


class Row(dict):
    """dict with attribute access, like ``pyspark.sql.Row``."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __getitem__(self, k):
        if isinstance(k, int):
            return list(self.values())[k]
        return dict.__getitem__(self, k)

    def asDict(self) -> dict:
        return dict(self)


class _NaFunctions:
    def __init__(self, df: 'DataFrame'):
        self._df = df

    def drop(self) -> 'DataFrame':
        keep = np.ones(self._df.count(), dtype=bool)
        for c in self._df._cols.values():
            if np.issubdtype(c.dtype, np.floating):
                nan = np.isnan(c)
                if nan.ndim > 1:
                    nan = nan.reshape(nan.shape[0], -1).any(axis=1)
                keep &= ~nan
        return self._df._take(np.nonzero(keep)[0])


class DataFrame:
    def __init__(self, columns: Dict[str, np.ndarray], num_partitions: int = 1):
        n = None
        self._cols: Dict[str, np.ndarray] = {}
        for k, v in columns.items():
            a = np.asarray(v)
            if n is None:
                n = a.shape[0]
            elif a.shape[0] != n:
                raise ValueError(f'column {k!r} has {a.shape[0]} rows, expected {n}')
            self._cols[k] = a
        self._n = 0 if n is None else int(n)
        self.num_partitions = max(1, int(num_partitions))
        self._cached = False

    # ------------------------------------------------------------ constructors
    @staticmethod
    def from_pandas(pdf, num_partitions: int = 1) -> 'DataFrame':
        return DataFrame({c: pdf[c].to_numpy() for c in pdf.columns}, num_partitions)

    @staticmethod
    def from_arrow(table, num_partitions: int = 1) -> 'DataFrame':
        cols = {}
        for name in table.column_names:
            col = table.column(name)
            try:
                cols[name] = col.to_numpy(zero_copy_only=False)
            except Exception:
                cols[name] = np.asarray(col.to_pylist())
            if cols[name].dtype == object:  # list<...> columns
                cols[name] = np.asarray(col.to_pylist())
        return DataFrame(cols, num_partitions)

    @staticmethod
    def from_spark(sdf) -> 'DataFrame':  # pragma: no cover - needs pyspark
        """Collect a real ``pyspark.sql.DataFrame`` through Arrow."""
        pdf = sdf.toPandas()
        cols = {}
        for c in pdf.columns:
            v = pdf[c].to_numpy()
            if v.dtype == object and len(v) and hasattr(v[0], 'toArray'):
                v = np.stack([x.toArray() for x in v])
            elif v.dtype == object and len(v) and isinstance(v[0], (list, tuple, np.ndarray)):
                v = np.asarray([np.asarray(x) for x in v])
            cols[c] = v
        return DataFrame(cols, sdf.rdd.getNumPartitions())

    @staticmethod
    def read_csv(path: str, header: bool = True, num_partitions: int = 1) -> 'DataFrame':
        """CSV with header + schema inference (spark-csv ``header/inferSchema`` options,
        /root/reference/ddl_mnist_aztk.py:100-107)."""
        import pandas as pd
        pdf = pd.read_csv(path, header=0 if header else None)
        return DataFrame.from_pandas(pdf, num_partitions)

    # ---------------------------------------------------------------- basics
    @property
    def columns(self) -> List[str]:
        return list(self._cols.keys())

    def __getitem__(self, name: str) -> np.ndarray:
        return self._cols[name]

    def __contains__(self, name: str) -> bool:
        return name in self._cols

    def count(self) -> int:
        return self._n

    def __len__(self) -> int:
        return self._n

    @property
    def na(self) -> _NaFunctions:
        return _NaFunctions(self)

    @property
    def schema(self) -> Dict[str, str]:
        out = {}
        for k, v in self._cols.items():
            if v.ndim == 1:
                out[k] = str(v.dtype)
            elif v.ndim == 2:
                out[k] = 'vector'
            else:
                out[k] = 'array<' * (v.ndim - 1) + 'double' + '>' * (v.ndim - 1)
        return out

    def printSchema(self, print_fn=print) -> None:
        print_fn('root')
        for k, t in self.schema.items():
            print_fn(f' |-- {k}: {t}')

    def _take(self, idx: np.ndarray) -> 'DataFrame':
        return DataFrame({k: v[idx] for k, v in self._cols.items()}, self.num_partitions)

    def select(self, *cols) -> 'DataFrame':
        if len(cols) == 1 and isinstance(cols[0], (list, tuple)):
            cols = tuple(cols[0])
        return DataFrame({c: self._cols[c] for c in cols}, self.num_partitions)

    def drop(self, *cols) -> 'DataFrame':
        return DataFrame({k: v for k, v in self._cols.items() if k not in cols}, self.num_partitions)

    def withColumn(self, name: str, values: Union[np.ndarray, Callable[['DataFrame'], np.ndarray]]) -> 'DataFrame':
        v = values(self) if callable(values) else values
        cols = dict(self._cols)
        cols[name] = np.asarray(v)
        return DataFrame(cols, self.num_partitions)

    def withColumnRenamed(self, old: str, new: str) -> 'DataFrame':
        return DataFrame({(new if k == old else k): v for k, v in self._cols.items()}, self.num_partitions)

    def limit(self, n: int) -> 'DataFrame':
        return self._take(np.arange(min(max(int(n), 0), self._n)))

    def orderBy(self, col: str, ascending: bool = True) -> 'DataFrame':
        idx = np.argsort(self._cols[col], kind='stable')
        if not ascending:
            idx = idx[::-1]
        return self._take(idx)

    sort = orderBy

    def where(self, mask: np.ndarray) -> 'DataFrame':
        return self._take(np.nonzero(np.asarray(mask, dtype=bool))[0])

    filter = where

    def shuffle(self, seed: Optional[int] = None) -> 'DataFrame':
        """``distkeras.utils.shuffle`` analogue (orderBy(rand()))."""
        rng = np.random.default_rng(seed)
        return self._take(rng.permutation(self._n))

    def repartition(self, n: int) -> 'DataFrame':
        return DataFrame(self._cols, n)

    coalesce = repartition

    def cache(self) -> 'DataFrame':
        self._cached = True
        return self

    def union(self, other: 'DataFrame') -> 'DataFrame':
        return DataFrame({k: np.concatenate([v, other._cols[k]]) for k, v in self._cols.items()}, self.num_partitions)

    # ------------------------------------------------------------ partitions
    def partition_indices(self, i: int, n: Optional[int] = None) -> np.ndarray:
        """Row indices of partition ``i`` (contiguous blocks, sizes differ by at most one)."""
        n = self.num_partitions if n is None else n
        base, rem = divmod(self._n, n)
        start = i * base + min(i, rem)
        return np.arange(start, start + base + (1 if i < rem else 0))

    def partition(self, i: int, n: Optional[int] = None) -> 'DataFrame':
        p = self._take(self.partition_indices(i, n))
        p.num_partitions = 1
        return p

    def partitions(self, n: Optional[int] = None) -> List['DataFrame']:
        n = self.num_partitions if n is None else n
        return [self.partition(i, n) for i in range(n)]

    def mapPartitionsWithIndex(self, fn: Callable[[int, 'DataFrame'], Iterable]) -> List:
        out = []
        for i, p in enumerate(self.partitions()):
            out.extend(list(fn(i, p)))
        return out

    # --------------------------------------------------------------- actions
    def collect(self) -> List[Row]:
        keys = self.columns
        cols = [self._cols[k] for k in keys]
        return [Row({k: (c[i].tolist() if c.ndim > 1 else c[i].item() if hasattr(c[i], 'item') else c[i])
                     for k, c in zip(keys, cols)}) for i in range(self._n)]

    def first(self) -> Row:
        return self.limit(1).collect()[0]

    def toPandas(self):
        import pandas as pd
        return pd.DataFrame({k: (list(v) if v.ndim > 1 else v) for k, v in self._cols.items()})

    def agg_min(self, col: str):
        return self._cols[col].min().item()

    def agg_max(self, col: str):
        return self._cols[col].max().item()

    def show(self, n: int = 20, truncate: int = 20, print_fn=print) -> None:
        keys = self.columns
        rows = []
        for i in range(min(n, self._n)):
            r = []
            for k in keys:
                v = self._cols[k][i]
                s = np.array2string(np.asarray(v), separator=',', threshold=6) if np.ndim(v) else str(v)
                s = s.replace('\n', '')
                if truncate and len(s) > truncate:
                    s = s[:truncate - 3] + '...'
                r.append(s)
            rows.append(r)
        widths = [max(len(k), *(len(r[j]) for r in rows)) if rows else len(k) for j, k in enumerate(keys)]
        sep = '+' + '+'.join('-' * w for w in widths) + '+'
        print_fn(sep)
        print_fn('|' + '|'.join(k.rjust(w) for k, w in zip(keys, widths)) + '|')
        print_fn(sep)
        for r in rows:
            print_fn('|' + '|'.join(c.rjust(w) for c, w in zip(r, widths)) + '|')
        print_fn(sep)
        if self._n > n:
            print_fn(f'only showing top {n} row' + ('s' if n != 1 else ''))


    def to_string(self, n: int = 20, truncate: int = 20) -> str:
        """What ``show`` prints, as a string (run records, logs)."""
        lines: List[str] = []
        self.show(n, truncate, print_fn=lines.append)
        return '\n'.join(lines)


class VectorAssembler:
    """``pyspark.ml.feature.VectorAssembler`` analogue: stack scalar/vector columns into one vector
    column (/root/reference/ddl_mnist_aztk.py:118, /root/reference/ddl_nyiso_aztk.py:154,165,169)."""

    def __init__(self, inputCols: Sequence[str], outputCol: str):
        self.inputCols = list(inputCols)
        self.outputCol = outputCol

    def transform(self, df: DataFrame) -> DataFrame:
        parts = []
        for c in self.inputCols:
            a = np.asarray(df[c], dtype=np.float64)
            parts.append(a.reshape(a.shape[0], -1))
        return df.withColumn(self.outputCol, np.concatenate(parts, axis=1))
