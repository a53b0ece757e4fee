"""Synthetic stand-ins for the reference's two datasets (there is no network or blob access).

* MNIST-shaped: a ``label`` column plus 784 ``pixelN`` uint8 columns, like the CSV read at
  /root/reference/ddl_mnist_aztk.py:100-107.  Each class draws a distinct blob pattern so a
  model can actually learn it.
* NYISO-shaped: hourly ``TimeStamp, Name, HourAvgLoad, temperature`` rows starting
  2016-01-02 like /root/reference/ddl_nyiso_hdi.ipynb:207-209, with daily/weekly/annual
  seasonality so the forecasters converge.
"""
from __future__ import annotations

import numpy as np

from .frame import DataFrame


def mnist_like_arrays(n: int, seed: int = 0, nb_classes: int = 10):
    """Returns (pixels uint8 [n,784], labels int64 [n])."""
    rng = np.random.default_rng(seed)
    proto_rng = np.random.default_rng(12345)  # class prototypes are fixed across seeds/ranks
    yy, xx = np.mgrid[0:28, 0:28]
    protos = np.zeros((nb_classes, 28, 28), dtype=np.float32)
    for c in range(nb_classes):
        for _ in range(3):
            cy, cx = proto_rng.uniform(6, 22, size=2)
            s = proto_rng.uniform(2.0, 4.0)
            protos[c] += np.exp(-((yy - cy) ** 2 + (xx - cx) ** 2) / (2 * s * s))
        protos[c] /= protos[c].max()
    labels = rng.integers(0, nb_classes, size=n)
    img = protos[labels] * rng.uniform(0.6, 1.0, size=(n, 1, 1)).astype(np.float32)
    img = img + rng.normal(0, 0.08, size=img.shape).astype(np.float32)
    pixels = np.clip(img * 255.0, 0, 255).astype(np.uint8).reshape(n, 784)
    return pixels, labels.astype(np.int64)


def mnist_like_frame(n: int, seed: int = 0, as_columns: bool = False) -> DataFrame:
    pixels, labels = mnist_like_arrays(n, seed)
    if as_columns:  # exactly the CSV's shape: 785 scalar columns
        cols = {'label': labels}
        for i in range(784):
            cols[f'pixel{i}'] = pixels[:, i]
        return DataFrame(cols)
    return DataFrame({'label': labels, 'features': pixels.astype(np.float64)})


def nyiso_like_frame(n_hours: int = 11688, seed: int = 0, name: str = 'N.Y.C.') -> DataFrame:
    """Hourly load/temperature series; default length gives 11,519 windowed training rows + 120 test
    rows + edge rows, the count implied by ddl_nyiso_hdi.ipynb:608."""
    rng = np.random.default_rng(seed)
    t = np.arange(n_hours, dtype=np.float64)
    ts = np.datetime64('2016-01-02T00:00:00') + (t.astype(np.int64) * np.timedelta64(3600, 's'))
    day = 2 * np.pi * (t % 24) / 24.0
    week = 2 * np.pi * (t % 168) / 168.0
    year = 2 * np.pi * t / (24 * 365.25)
    temp = 55.0 - 25.0 * np.cos(year) + 6.0 * np.sin(day - 2.0) + rng.normal(0, 1.5, n_hours)
    load = (5800.0 + 1200.0 * np.sin(day - 1.8) + 350.0 * np.sin(2 * day) - 250.0 * np.cos(week)
            + 18.0 * np.abs(temp - 60.0) + rng.normal(0, 60.0, n_hours))
    return DataFrame({'TimeStamp': ts, 'Name': np.array([name] * n_hours), 'HourAvgLoad': load,
                      'temperature': temp})


def nyiso_windows_arrays(n: int, seq_len: int = 25, seed: int = 0):
    """Already-windowed synthetic tensors for benchmarks: x [n,seq_len,1] float32 in [0,1], y [n,1]."""
    rng = np.random.default_rng(seed)
    phase = rng.uniform(0, 2 * np.pi, size=(n, 1))
    t = np.arange(seq_len + 1)[None, :]
    s = 0.5 + 0.35 * np.sin(phase + 2 * np.pi * t / 24.0) + rng.normal(0, 0.02, size=(n, seq_len + 1))
    x = s[:, :seq_len, None].astype(np.float32)
    y = s[:, seq_len:seq_len + 1].astype(np.float32)
    return x, y
