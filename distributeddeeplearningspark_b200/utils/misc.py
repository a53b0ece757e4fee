"""Small helpers mirroring ``distkeras.utils`` pieces the scripts touch
(``get_os_username`` at /root/reference/ddl_mnist_aztk.py:71) plus env knobs."""
from __future__ import annotations

import getpass
import os
import socket


def get_os_username() -> str:
    try:
        return getpass.getuser()
    except Exception:  # no passwd entry in minimal containers
        return os.environ.get('USER', 'unknown')


def determine_host_address() -> str:
    """Address other ranks can reach this host at; loopback when the hostname does not resolve."""
    try:
        return socket.gethostbyname(socket.gethostname())
    except OSError:
        return '127.0.0.1'


def free_port(host: str = '127.0.0.1') -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind((host, 0))
        return s.getsockname()[1]


def env_int(name: str, default: int) -> int:
    v = os.environ.get(name)
    return default if v in (None, '') else int(v)


def env_str(name: str, default: str) -> str:
    v = os.environ.get(name)
    return default if v in (None, '') else v


def env_flag(name: str, default: bool = False) -> bool:
    v = os.environ.get(name)
    if v in (None, ''):
        return default
    return v.lower() not in ('0', 'false', 'no', 'off')


# ---------------------------------------------------------------------------------------------
# The rest of the ``distkeras.utils`` surface.  Imports are local: these
# helpers sit above the data / model layers and this module is imported by both.
def serialize_keras_model(model) -> dict:
    """``{'model': <architecture JSON>, 'weights': [ndarray, ...]}`` — the blob dist-keras ships to workers."""
    from ..models.sequential import serialize_model
    return serialize_model(model)


def deserialize_keras_model(blob: dict):
    from ..models.sequential import deserialize_model
    return deserialize_model(blob)


def new_dataframe_row(old_row, column_name: str, column_value):
    """Copy of ``old_row`` with one more (or replaced) column, like ``distkeras.utils.new_dataframe_row``."""
    from ..data.frame import Row
    d = old_row.asDict() if hasattr(old_row, 'asDict') else dict(old_row)
    d[column_name] = column_value
    return Row(**d)


def shuffle(dataframe, seed=None):
    """Random row order (dist-keras: ``orderBy(rand())``)."""
    return dataframe.shuffle(seed)


def precache(dataframe, num_workers: int = None):
    """Materialise (and optionally repartition) a frame before training; returns the row count like the
    ``count()`` dist-keras uses to force the cache."""
    if num_workers:
        dataframe = dataframe.repartition(num_workers)
    dataframe.cache()
    return dataframe.count()


def to_one_hot_encoded_dense(value, n_dim: int = 2):
    import numpy as np
    v = np.zeros(int(n_dim), dtype=np.float64)
    v[int(value)] = 1.0
    return v


def history_executors_average(history):
    """Average per-worker training histories position by position (shortest history wins)."""
    import numpy as np
    history = [h for h in history if len(h)]
    if not history:
        return []
    n = min(len(h) for h in history)
    return np.mean(np.asarray([h[:n] for h in history], dtype=np.float64), axis=0).tolist()


def history_executor(history, worker_id: int):
    return list(history[worker_id]) if 0 <= worker_id < len(history) else []
