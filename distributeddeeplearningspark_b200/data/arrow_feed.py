"""Arrow record batches -> pinned host ring -> async H2D.

Spark hands a Python worker its partition as a stream of Arrow record batches (``DataFrame.mapInArrow``, the
columnar replacement of the row-at-a-time pickling every dist-keras transformer and worker pays,
SURVEY.md §3.1 / §2.4 "partitions -> device").  ``ArrowStreamFeeder`` consumes such a stream: the fixed-size-list
columns of each record batch are viewed as NumPy blocks without a copy, rows are cut into minibatches of
``batch_size`` straight into the slots of a pinned ring buffer (a minibatch may straddle record batches and epoch
boundaries, like dist-keras' batching), and each full slot is copied to the device with ``cudaMemcpyAsync`` on a copy
stream, ``depth`` minibatches ahead of the compute stream.

Without a JVM the same path is driven from Arrow IPC files: the driver writes one file per partition
(``write_partition_file``), ranks memory-map theirs and stream its record batches (``ipc_file_source``) — that is what
the trainers do with ``DDLS_PARTITION_FORMAT=arrow``.
"""
from __future__ import annotations

from typing import Callable, Dict, Iterable, Iterator, Optional, Sequence, Tuple

import numpy as np
import torch

try:  # pyarrow ships with the image; the module degrades to an ImportError at use, not at import of the package
    import pyarrow as pa
    import pyarrow.ipc as pa_ipc
except ImportError:  # pragma: no cover
    pa = None
    pa_ipc = None


def _need_arrow():
    if pa is None:  # pragma: no cover
        raise ImportError('pyarrow is required for the Arrow partition feed')


def column_to_numpy(col, row_shape: Tuple[int, ...]) -> np.ndarray:
    """Rows of an Arrow column as a NumPy block ``[rows, *row_shape]`` — zero-copy for fixed-size lists of primitives and
    for primitive columns; nested Python lists (``array<array<double>>`` as in ipynb:516) are flattened with a copy."""
    if isinstance(col, pa.ChunkedArray):
        col = col.combine_chunks()
    n = len(col)
    flat = col
    while pa.types.is_fixed_size_list(flat.type) or pa.types.is_list(flat.type) or pa.types.is_large_list(flat.type):
        flat = flat.flatten()
    try:
        a = flat.to_numpy(zero_copy_only=True)
    except pa.ArrowInvalid:
        a = flat.to_numpy(zero_copy_only=False)
    return a.reshape((n,) + tuple(row_shape))


def write_partition_file(path: str, columns: Dict[str, np.ndarray], rows_per_batch: int = 4096) -> int:
    """One Arrow IPC file for a partition: every column becomes a fixed-size list of its row's elements (dtype kept:
    uint8 pixels stay uint8).  Returns the number of record batches written."""
    _need_arrow()
    n = len(next(iter(columns.values())))
    fields, shapes = [], {}
    for k, v in columns.items():
        v = np.ascontiguousarray(v)
        width = int(np.prod(v.shape[1:])) if v.ndim > 1 else 1
        fields.append(pa.field(k, pa.list_(pa.from_numpy_dtype(v.dtype), width),
                               metadata={'row_shape': ','.join(str(s) for s in v.shape[1:])}))
        shapes[k] = (v, width)
    schema = pa.schema(fields)
    nb = 0
    with pa_ipc.new_file(path, schema) as w:
        for lo in range(0, n, rows_per_batch):
            hi = min(n, lo + rows_per_batch)
            arrays = [pa.FixedSizeListArray.from_arrays(pa.array(v[lo:hi].reshape(-1)), width) for v, width in shapes.values()]
            w.write_batch(pa.record_batch(arrays, schema=schema))
            nb += 1
    return nb


def ipc_file_source(path: str) -> Tuple[Callable[[], Iterator], int, Dict[str, Tuple[Tuple[int, ...], np.dtype]]]:
    """(source, rows, {column: (row_shape, dtype)}) for a partition file: ``source()`` yields its memory-mapped record
    batches (one pass = one epoch)."""
    _need_arrow()
    reader = pa_ipc.open_file(pa.memory_map(path, 'r'))
    rows = sum(reader.get_batch(i).num_rows for i in range(reader.num_record_batches))
    spec = {}
    for f in reader.schema:
        meta = (f.metadata or {}).get(b'row_shape', b'').decode()
        shape = tuple(int(s) for s in meta.split(',') if s) if meta else (f.type.list_size,)
        spec[f.name] = (shape, np.dtype(f.type.value_type.to_pandas_dtype()))

    def source():
        for i in range(reader.num_record_batches):
            yield reader.get_batch(i)
    return source, rows, spec


class ArrowStreamFeeder:
    """Minibatches from a stream of Arrow record batches.

    ``source``: a callable returning an iterator of ``pyarrow.RecordBatch`` (called once per epoch) — an IPC file
    (``ipc_file_source``) or, inside a Spark barrier stage, the iterator ``mapInArrow`` passes to the worker (one
    epoch, unless the stage caches the batches).  ``columns``: ``{output name: (arrow column, row_shape, dtype)}``.
    Iteration yields dicts of device tensors ``[batch_size, *row_shape]``; the tensors of a slot are reused after
    ``depth`` minibatches."""

    def __init__(self, source: Callable[[], Iterable], columns: Dict[str, Tuple[str, Sequence[int], np.dtype]],
                 batch_size: int, device, num_epoch: int = 1, depth: int = 3, total_rows: Optional[int] = None):
        _need_arrow()
        self.source, self.columns = source, {k: (c, tuple(s), np.dtype(d)) for k, (c, s, d) in columns.items()}
        self.batch_size, self.num_epoch = int(batch_size), int(num_epoch)
        self.device = torch.device(device)
        self.cuda = self.device.type == 'cuda'
        self.depth = max(2, int(depth))
        self.total_rows = total_rows
        mk = lambda s, d, **kw: torch.empty((self.batch_size,) + s, dtype=torch.from_numpy(np.empty(0, dtype=d)).dtype, **kw)
        self._host = [{k: mk(s, d, pin_memory=self.cuda) for k, (_, s, d) in self.columns.items()} for _ in range(self.depth)]
        self.h2d_bytes_per_batch = sum(t.numel() * t.element_size() for t in self._host[0].values())
        if self.cuda:
            self._dev = [{k: torch.empty_like(t, device=self.device) for k, t in slot.items()} for slot in self._host]
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._ready = [torch.cuda.Event() for _ in range(self.depth)]
            self._free = [torch.cuda.Event() for _ in range(self.depth)]

    def __len__(self) -> int:
        if self.total_rows is None:
            raise TypeError('length unknown: pass total_rows (a mapInArrow stream does not announce its size)')
        return (self.total_rows * self.num_epoch) // self.batch_size

    def _filled_slots(self) -> Iterator[int]:
        """Cut the row stream into pinned slots; yields a slot index each time one is full and on its way to the device."""
        slot, filled, used = 0, 0, [False] * self.depth
        for _ in range(self.num_epoch):
            for rb in self.source():
                blocks = {k: column_to_numpy(rb.column(rb.schema.get_field_index(c)), s) for k, (c, s, _) in self.columns.items()}
                n, off = rb.num_rows, 0
                while off < n:
                    if filled == 0 and self.cuda and used[slot]:
                        self._free[slot].synchronize()          # compute finished with this slot's device tensors
                    take = min(self.batch_size - filled, n - off)
                    for k, blk in blocks.items():
                        self._host[slot][k][filled:filled + take].copy_(torch.from_numpy(blk[off:off + take]))
                    filled += take; off += take
                    if filled == self.batch_size:
                        if self.cuda:
                            with torch.cuda.stream(self._copy_stream):
                                for k, t in self._host[slot].items():
                                    self._dev[slot][k].copy_(t, non_blocking=True)
                                self._ready[slot].record(self._copy_stream)
                        used[slot] = True
                        yield slot
                        slot, filled = (slot + 1) % self.depth, 0
        # a final partial minibatch is dropped, as dist-keras does

    def __iter__(self) -> Iterator[Dict[str, torch.Tensor]]:
        it = self._filled_slots()
        pending = []
        for _ in range(self.depth - 1):                          # run ahead of the consumer
            s = next(it, None)
            if s is None:
                break
            pending.append(s)
        while pending:
            cur = pending.pop(0)
            nxt = next(it, None)
            if nxt is not None:
                pending.append(nxt)
            if self.cuda:
                torch.cuda.current_stream(self.device).wait_event(self._ready[cur])
                yield self._dev[cur]
                self._free[cur].record(torch.cuda.current_stream(self.device))
            else:
                yield {k: t.clone() for k, t in self._host[cur].items()}
