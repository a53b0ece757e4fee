"""Trainer facades with the dist-keras constructor/usage shape the reference scripts use.

    trainer = ADAG(keras_model=model, worker_optimizer='adam', loss='categorical_crossentropy',
                   num_workers=8, batch_size=16, communication_window=5, num_epoch=5,
                   features_col='matrix', label_col='label_encoded')
    trained = trainer.train(training_set)
    trainer.get_training_time(); trainer.parameter_server.num_updates

(/root/reference/ddl_mnist_aztk.py:216-224, /root/reference/ddl_nyiso_aztk.py:207-218,255-261;
``DynSGD`` at ddl_nyiso_aztk.py:212-215; the "SingleWorker" comparison at ddl_mnist_aztk.py:215.)

``train`` fans out through ``spark.run`` — one rank per partition/GPU — and each
rank runs the fused step engine over its partition.  ``AllreduceTrainer`` is the
synchronous Horovod-style path (``DistributedOptimizer``) with the same facade.
"""
from __future__ import annotations

import os
import shutil
import tempfile
import time
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .. import runtime
from ..data.feeder import DeviceFeeder
from ..data.frame import DataFrame
from ..models.sequential import Sequential, deserialize_model, serialize_model
from ..utils.logging import MetricsWriter, get_logger
from ..utils.timing import PhaseTimers, Timeline
from .engine import StepEngine
from .parameter_server import CollectiveCenter, MulticastCenter, ParameterServerView


# ----------------------------------------------------------------------------
# partition staging (driver -> ranks)
# ----------------------------------------------------------------------------
def _feature_dtype(a: np.ndarray) -> np.dtype:
    return np.uint8 if a.dtype == np.uint8 else np.float32


def partition_format() -> str:
    """'npy' (memory-mapped arrays, whole-partition random access: needed for shuffling) or 'arrow' (Arrow IPC record
    batches streamed through ``data.arrow_feed.ArrowStreamFeeder`` — the form a Spark ``mapInArrow`` stage delivers)."""
    return os.environ.get('DDLS_PARTITION_FORMAT', 'npy').lower()


def stage_partitions(df: DataFrame, features_col: str, label_col: str, n: int, root: Optional[str] = None) -> str:
    """Write partition i's feature/label columns where the ranks can map them (driver side): ``x{i}.npy`` / ``y{i}.npy``,
    or one Arrow IPC file ``part{i}.arrow`` per partition with ``DDLS_PARTITION_FORMAT=arrow``."""
    base = root or ('/dev/shm' if os.path.isdir('/dev/shm') else None)
    d = tempfile.mkdtemp(prefix='ddls_parts_', dir=base)
    arrow = partition_format() == 'arrow'
    for i in range(n):
        idx = df.partition_indices(i, n)
        x = np.asarray(df[features_col])[idx]
        y = np.asarray(df[label_col])[idx]
        x = np.ascontiguousarray(x, dtype=_feature_dtype(x)); y = np.ascontiguousarray(y, dtype=np.float32)
        if arrow:
            from ..data.arrow_feed import write_partition_file
            write_partition_file(os.path.join(d, f'part{i}.arrow'), {'x': x, 'y': y.reshape(len(y), -1)},
                                 rows_per_batch=int(os.environ.get('DDLS_ARROW_BATCH_ROWS', '2048')))
        else:
            np.save(os.path.join(d, f'x{i}.npy'), x)
            np.save(os.path.join(d, f'y{i}.npy'), y)
    return d


def load_partition(d: str, i: int):
    return np.load(os.path.join(d, f'x{i}.npy'), mmap_mode='r'), np.load(os.path.join(d, f'y{i}.npy'), mmap_mode='r')


def open_feeder(spec: Dict, rank: int, dev):
    """The rank's minibatch feeder over its staged partition (pinned ring + copy stream in both formats)."""
    arrow_path = os.path.join(spec['parts'], f'part{rank}.arrow')
    if os.path.exists(arrow_path):
        from ..data.arrow_feed import ArrowStreamFeeder, ipc_file_source
        if spec.get('shuffle'):
            raise ValueError('shuffle=True needs random access: stage partitions with DDLS_PARTITION_FORMAT=npy')
        source, rows, cols = ipc_file_source(arrow_path)
        return ArrowStreamFeeder(source, {k: (k, shape, dt) for k, (shape, dt) in cols.items()}, spec['batch_size'], dev,
                                 num_epoch=spec['num_epoch'], total_rows=rows)
    x, y = load_partition(spec['parts'], rank)
    return DeviceFeeder({'x': np.asarray(x), 'y': np.asarray(y)}, spec['batch_size'], dev,
                        num_epoch=spec['num_epoch'], shuffle=spec.get('shuffle', False),
                        seed=spec.get('seed') or 0, carry_over=True)


# ----------------------------------------------------------------------------
# worker bodies (run inside each rank)
# ----------------------------------------------------------------------------
def _worker(spec: Dict) -> Dict:
    # Observability (the reference only has get_training_time()): host-side phase timers are always on and
    # come back with the result; DDLS_TIMELINE=<path> writes a Chrome trace per rank (Horovod-timeline analogue);
    # DDLS_METRICS=<path> appends {step, loss, samples_per_s} every DDLS_METRICS_EVERY steps (rank 0) as JSON lines.
    phases = PhaseTimers()
    t_init = time.perf_counter()
    runtime.init()
    rank, size = runtime.rank(), runtime.size()
    dev = runtime.device()
    comm = runtime.communicator()
    timeline = Timeline(rank=rank)
    metrics = MetricsWriter(os.environ.get('DDLS_METRICS'), rank=rank)
    metrics_every = max(1, int(os.environ.get('DDLS_METRICS_EVERY', '50')))
    phases.total['init'] = time.perf_counter() - t_init; phases.count['init'] = 1
    timeline.instant('runtime.init done', size=size, comm=type(comm).__name__)
    if spec.get('seed') is not None:
        torch.manual_seed(spec['seed'] + rank)
    with phases.phase('load'), timeline.span('load model + partition', cat='setup'):
        model = deserialize_model(spec['model']).to(dev)
        feeder = open_feeder(spec, rank, dev)
    mode = spec['mode']
    engine = StepEngine(model, optimizer=spec['worker_optimizer'], loss=spec['loss'], device=dev,
                        comm=comm if mode == 'allreduce' else None, distributed=(mode == 'allreduce'),
                        optimizer_kwargs=spec.get('optimizer_kwargs'))
    flat = engine.optimizer.flat if hasattr(engine.optimizer, 'flat') else engine.optimizer.inner.flat
    if size > 1:
        comm.broadcast_(flat.flat_params, 0)
    window = int(spec.get('communication_window', 1))
    center = None
    async_center = False
    if mode in ('adag', 'dynsgd', 'downpour', 'aeasgd'):
        # GPUs with a peer-memory communicator: asynchronous center in multicast memory (DDLS_ADAG=sync forces
        # the lock-step collective emulation, which is also what CPU/gloo and single-rank runs use)
        async_center = (type(comm).__name__ == 'PeerMemoryComm' and os.environ.get('DDLS_ADAG', 'async') != 'sync'
                        and mode != 'aeasgd')       # the elastic exchange is a lock-step collective
        cls = MulticastCenter if async_center else CollectiveCenter
        center = cls(comm, flat.flat_params, staleness_aware=(mode == 'dynsgd'))
    n_batches = len(feeder)
    center_modes = ('adag', 'dynsgd', 'downpour', 'aeasgd')
    avg_every = 0
    if mode in center_modes:
        my_rounds = n_batches // window
        max_rounds = max(comm.allgather_object(my_rounds))
    elif mode in ('allreduce', 'averaging'):
        # lock-step collectives: every rank must run the same number of steps
        n_batches = min(comm.allgather_object(n_batches))
        if mode == 'averaging':                        # weights are averaged at every epoch boundary
            avg_every = max(1, n_batches // max(1, int(spec['num_epoch'])))
    alpha = float(spec.get('rho', 5.0)) * float(spec.get('learning_rate', 0.1))
    losses: List[torch.Tensor] = []
    t0 = time.perf_counter()
    it = 0
    feed = iter(feeder)
    while it < n_batches:
        with phases.phase('feed'):
            batch = next(feed, None)
        if batch is None:
            break
        xb = batch['x']
        if xb.dtype != torch.float32:
            xb = xb.float()
        # step() returns the engine's static loss tensor on every CUDA-graph replay: snapshot it (a tiny on-stream copy),
        # otherwise the history would alias one buffer and read [.., last, last, last]
        with phases.phase('step'), timeline.span('step', step=it):
            losses.append(engine.step(xb, batch['y']).detach().clone())
        it += 1
        if center is not None and it % window == 0:
            with phases.phase('exchange'), timeline.span('commit+pull', cat='comm', update=it // window):
                if mode == 'aeasgd':
                    center.elastic_round(flat.flat_params, alpha, active=True)
                    engine.notify_params_changed()
                else:
                    center.commit_pull(flat.flat_params, 1.0 / window if mode == 'adag' else 1.0, active=True)
        if metrics.enabled and it % metrics_every == 0:
            metrics.write(step=it, loss=float(losses[-1]), samples_per_s=it * spec['batch_size'] / (time.perf_counter() - t0),
                          mode=mode)
        if avg_every and it % avg_every == 0 and size > 1:
            from ..parallel.comm import Average
            comm.allreduce_(flat.flat_params, Average)
            engine.notify_params_changed()
    if center is not None:
        if async_center:
            center.finalize()
        else:
            for _ in range(max_rounds - my_rounds):  # keep peers' collectives matched
                if mode == 'aeasgd':
                    center.elastic_round(flat.flat_params, alpha, active=False)
                else:
                    center.commit_pull(flat.flat_params, 0.0, active=False)
        flat.flat_params.copy_(center.center)
        num_updates = center.total_updates()
    else:
        num_updates = it * (size if mode == 'allreduce' else 1)
    if dev.type == 'cuda':
        torch.cuda.synchronize(dev)
    elapsed = time.perf_counter() - t0
    hist = torch.stack(losses).float().cpu().tolist() if losses else []
    out = {'rank': rank, 'history': hist, 'num_updates': num_updates, 'elapsed': elapsed, 'steps': it,
           'phases': phases.summary()}
    metrics.write(step=it, loss=(hist[-1] if hist else None), samples_per_s=it * spec['batch_size'] / max(elapsed, 1e-9),
                  mode=mode, final=True)
    metrics.close()
    timeline.dump()
    if rank == 0 or mode == 'ensemble':
        out['weights'] = model.get_weights()
        if spec.get('checkpoint'):
            from ..ckpt import save_checkpoint
            save_checkpoint(spec['checkpoint'], model, engine.optimizer)
    return out


# ----------------------------------------------------------------------------
# facades
# ----------------------------------------------------------------------------
class Trainer:
    """Base: holds the serialised master model and the training hyper-parameters."""

    mode = 'single'

    def __init__(self, keras_model: Sequential, loss: str, worker_optimizer: str = 'adam', metrics=('accuracy',),
                 features_col: str = 'features', label_col: str = 'label', batch_size: int = 32, num_epoch: int = 1,
                 num_workers: int = 1, backend=None, seed: Optional[int] = None, checkpoint: Optional[str] = None,
                 optimizer_kwargs: Optional[dict] = None, master_port: int = 5000, parallelism_factor: int = 1):
        self.master_model = serialize_model(keras_model)
        # dist-keras attributes kept for API parity: there is no socket server to bind (the center variable lives in
        # symmetric / multicast GPU memory), and one rank owns one GPU, so these two are informational only.
        self.master_port = int(master_port)
        self.parallelism_factor = int(parallelism_factor)
        self.loss = loss
        self.worker_optimizer = worker_optimizer
        self.metrics = list(metrics)
        self.features_column = features_col
        self.label_column = label_col
        self.batch_size = int(batch_size)
        self.num_epoch = int(num_epoch)
        self.num_workers = int(num_workers)
        self.backend = backend
        self.seed = seed
        self.checkpoint = checkpoint
        self.optimizer_kwargs = optimizer_kwargs
        self.history: List[List[float]] = []
        self.training_time_start = 0.0
        self.training_time_end = 0.0
        self.parameter_server = ParameterServerView()
        self.worker_results: List[Dict] = []

    # dist-keras surface --------------------------------------------------
    def record_training_start(self):
        self.training_time_start = time.time()

    def record_training_end(self):
        self.training_time_end = time.time()

    def get_training_time(self) -> float:
        return self.training_time_end - self.training_time_start

    def get_loop_time(self) -> float:
        """Seconds the slowest rank spent in its training loop (first minibatch to last update, device drained) — what is
        left of ``get_training_time()`` (driver wall clock, as in the reference) without process start, CUDA context
        creation, heap exchange and partition staging."""
        return max(getattr(self, 'loop_times', [0.0]) or [0.0])

    def get_history(self):
        return self.history

    def get_phase_times(self):
        """Per-rank host-side phase timers of the last ``train()``: {'init' | 'load' | 'feed' | 'step' | 'exchange':
        {'total_s', 'count', 'mean_ms'}} (the breakdown ``get_training_time()`` of the reference does not give)."""
        return getattr(self, 'phase_times', [])

    def get_averaged_history(self):
        if not self.history:
            return []
        n = min(len(h) for h in self.history)
        return np.mean(np.asarray([h[:n] for h in self.history]), axis=0).tolist()

    def has_history(self) -> bool:
        return any(len(h) for h in self.history)

    # ---------------------------------------------------------------- train
    def _spec(self, parts_dir: str) -> Dict:
        return {'mode': self.mode, 'model': self.master_model, 'parts': parts_dir, 'loss': self.loss,
                'worker_optimizer': self.worker_optimizer, 'batch_size': self.batch_size,
                'num_epoch': self.num_epoch, 'seed': self.seed, 'checkpoint': self.checkpoint,
                'optimizer_kwargs': self.optimizer_kwargs}

    def train(self, dataframe: DataFrame, shuffle: bool = False) -> Sequential:
        from .. import spark
        n = self.num_workers
        if shuffle:
            dataframe = dataframe.shuffle(self.seed)
        dataframe = dataframe.repartition(n).cache()
        parts_dir = stage_partitions(dataframe, self.features_column, self.label_column, n)
        spec = self._spec(parts_dir)
        spec['shuffle'] = False
        backend = self.backend
        if backend is None and n == 1:
            backend = 'inprocess'
        self.record_training_start()
        try:
            results = spark.run(_worker, args=(spec,), num_proc=n, backend=backend)
        finally:
            self.record_training_end()
            shutil.rmtree(parts_dir, ignore_errors=True)
        self.worker_results = results
        self.history = [r['history'] for r in results]
        self.phase_times = [r.get('phases', {}) for r in results]
        self.loop_times = [float(r.get('elapsed', 0.0)) for r in results]
        model = deserialize_model(self.master_model)
        model.set_weights(results[0]['weights'])
        self.parameter_server = ParameterServerView(self._count_updates(results), model)
        get_logger().info('%s: %d worker(s), %d updates, %.3f s', type(self).__name__, n,
                          self.parameter_server.num_updates, self.get_training_time())
        return model

    def _count_updates(self, results) -> int:
        return int(results[0]['num_updates'])


class SingleTrainer(Trainer):
    """One worker, plain sequential minibatch descent (dist-keras ``SingleTrainer``)."""

    mode = 'single'

    def __init__(self, keras_model, worker_optimizer, loss, metrics=('accuracy',), features_col='features',
                 label_col='label', num_epoch=1, batch_size=32, **kw):
        kw.pop('num_workers', None)
        super().__init__(keras_model, loss, worker_optimizer, metrics, features_col, label_col, batch_size,
                         num_epoch, num_workers=1, **kw)


class DistributedTrainer(Trainer):
    def __init__(self, keras_model, worker_optimizer, loss, metrics=('accuracy',), num_workers=2, batch_size=32,
                 features_col='features', label_col='label', num_epoch=1, **kw):
        super().__init__(keras_model, loss, worker_optimizer, metrics, features_col, label_col, batch_size,
                         num_epoch, num_workers, **kw)


class ADAG(DistributedTrainer):
    """Accumulated-gradient normalisation: every ``communication_window`` local steps each worker
    commits ``(W - W_pulled) / communication_window`` to the center variable and pulls it."""

    mode = 'adag'

    def __init__(self, keras_model, worker_optimizer, loss, metrics=('accuracy',), num_workers=2, batch_size=32,
                 features_col='features', label_col='label', num_epoch=1, communication_window=12, **kw):
        super().__init__(keras_model, worker_optimizer, loss, metrics, num_workers, batch_size, features_col,
                         label_col, num_epoch, **kw)
        self.communication_window = int(communication_window)

    def _spec(self, parts_dir):
        s = super()._spec(parts_dir)
        s['communication_window'] = self.communication_window
        return s


class DynSGD(ADAG):
    """Staleness-aware variant: the center scales each residual by 1/(staleness + 1)."""

    mode = 'dynsgd'

    def __init__(self, keras_model, worker_optimizer, loss, metrics=('accuracy',), num_workers=2, batch_size=32,
                 features_col='features', label_col='label', num_epoch=1, communication_window=5, **kw):
        super().__init__(keras_model, worker_optimizer, loss, metrics, num_workers, batch_size, features_col,
                         label_col, num_epoch, communication_window, **kw)


class DOWNPOUR(ADAG):
    """dist-keras ``DOWNPOUR``: every ``communication_window`` steps a worker commits its accumulated weight change
    un-normalised (``center += W - W_pulled``) and pulls the center."""

    mode = 'downpour'

    def __init__(self, keras_model, worker_optimizer, loss, metrics=('accuracy',), num_workers=2, batch_size=32,
                 features_col='features', label_col='label', num_epoch=1, communication_window=5, **kw):
        super().__init__(keras_model, worker_optimizer, loss, metrics, num_workers, batch_size, features_col,
                         label_col, num_epoch, communication_window, **kw)


class AEASGD(ADAG):
    """dist-keras ``AEASGD`` (asynchronous elastic averaging SGD, run here as lock-step rounds): every
    ``communication_window`` steps worker and center are pulled towards each other by ``rho * learning_rate`` times
    their difference."""

    mode = 'aeasgd'

    def __init__(self, keras_model, worker_optimizer, loss, metrics=('accuracy',), num_workers=2, batch_size=32,
                 features_col='features', label_col='label', num_epoch=1, communication_window=32, rho=5.0,
                 learning_rate=0.1, **kw):
        super().__init__(keras_model, worker_optimizer, loss, metrics, num_workers, batch_size, features_col,
                         label_col, num_epoch, communication_window, **kw)
        self.rho, self.learning_rate = float(rho), float(learning_rate)

    def _spec(self, parts_dir):
        s = super()._spec(parts_dir)
        s['rho'], s['learning_rate'] = self.rho, self.learning_rate
        return s


class EAMSGD(AEASGD):
    """dist-keras ``EAMSGD``: elastic averaging with momentum SGD workers (``worker_optimizer='sgd'`` with
    ``momentum``, Nesterov by default as in the paper)."""

    def __init__(self, keras_model, worker_optimizer='sgd', loss='mean_squared_error', metrics=('accuracy',),
                 num_workers=2, batch_size=32, features_col='features', label_col='label', num_epoch=1,
                 communication_window=32, rho=5.0, learning_rate=0.1, momentum=0.9, **kw):
        okw = dict(kw.pop('optimizer_kwargs', None) or {})
        okw.setdefault('momentum', momentum)
        okw.setdefault('nesterov', True)
        super().__init__(keras_model, 'sgd', loss, metrics, num_workers, batch_size, features_col, label_col, num_epoch,
                         communication_window, rho, learning_rate, optimizer_kwargs=okw, **kw)
        self.momentum = float(momentum)


class AveragingTrainer(DistributedTrainer):
    """dist-keras ``AveragingTrainer``: workers train independently on their partitions and the weights are averaged
    at every epoch boundary (one allreduce of the flat parameter buffer per epoch)."""

    mode = 'averaging'

    def _count_updates(self, results) -> int:
        return int(self.num_epoch)


class EnsembleTrainer(DistributedTrainer):
    """dist-keras ``EnsembleTrainer``: ``num_models`` copies of the model are trained independently, one per
    partition, with no communication; ``train`` returns the list of trained models."""

    mode = 'ensemble'

    def __init__(self, keras_model, worker_optimizer, loss, metrics=('accuracy',), num_models=2, batch_size=32,
                 features_col='features', label_col='label', num_epoch=1, **kw):
        kw.pop('num_workers', None)
        super().__init__(keras_model, worker_optimizer, loss, metrics, num_models, batch_size, features_col,
                         label_col, num_epoch, **kw)
        self.num_models = int(num_models)

    def train(self, dataframe: DataFrame, shuffle: bool = False):
        super().train(dataframe, shuffle)
        models = []
        for r in self.worker_results:
            m = deserialize_model(self.master_model)
            m.set_weights(r['weights'])
            models.append(m)
        return models

    def _count_updates(self, results) -> int:
        return int(sum(r['steps'] for r in results))


class AllreduceTrainer(DistributedTrainer):
    """Synchronous data parallelism: ``DistributedOptimizer`` averages gradients every step
    (the Horovod-shaped product path of BASELINE.json)."""

    mode = 'allreduce'
