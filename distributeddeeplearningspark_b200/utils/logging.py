"""Rank-aware logger and a JSONL metrics sink.

The reference only ``print``s (/root/reference/ddl_mnist_aztk.py:221-224); here
rank 0 logs by default and per-step metrics can be appended to a JSONL file."""
from __future__ import annotations

import json
import logging
import os
import sys
import time
from typing import Optional

_FMT = '%(asctime)s %(name)s[r%(rank)s] %(levelname)s %(message)s'


class _RankFilter(logging.Filter):
    def filter(self, record):
        record.rank = os.environ.get('DDLS_RANK', os.environ.get('RANK', '0'))
        return True


def get_logger(name: str = 'ddls', all_ranks: bool = False) -> logging.Logger:
    log = logging.getLogger(name)
    if not log.handlers:
        h = logging.StreamHandler(sys.stderr)
        h.setFormatter(logging.Formatter(_FMT))
        h.addFilter(_RankFilter())
        log.addHandler(h)
        log.setLevel(os.environ.get('DDLS_LOG_LEVEL', 'INFO'))
        log.propagate = False
    rank = int(os.environ.get('DDLS_RANK', os.environ.get('RANK', '0')))
    if rank != 0 and not all_ranks:
        log.setLevel(logging.WARNING)
    return log


class MetricsWriter:
    """Append-only JSONL metrics file (rank 0 only unless ``all_ranks``)."""

    def __init__(self, path: Optional[str], rank: int = 0, all_ranks: bool = False):
        self.enabled = path is not None and (rank == 0 or all_ranks)
        self.rank = rank
        self._f = open(path, 'a') if self.enabled else None

    def write(self, **kv) -> None:
        if not self.enabled:
            return
        kv.setdefault('ts', time.time())
        kv.setdefault('rank', self.rank)
        self._f.write(json.dumps(kv) + '\n')
        self._f.flush()

    def close(self) -> None:
        if self._f is not None:
            self._f.close()
            self._f = None
