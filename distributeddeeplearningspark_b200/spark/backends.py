"""Barrier-stage backends for ``spark.run``.

The launcher is coded against a narrow protocol:

    backend.default_parallelism()                     -> int
    backend.run_barrier_stage(n, task)                -> [task(ctx) result for partition 0..n-1]
    ctx.partition_id() / ctx.all_gather(str) / ctx.barrier() / ctx.task_addresses()

``LocalBarrierBackend`` implements it with spawned processes and a coordinator
thread in the driver (tested on this image).  ``PySparkBarrierBackend`` maps it
onto ``sc.parallelize(range(n), n).barrier().mapPartitions(f).collect()`` and
``BarrierTaskContext`` — the barrier-mode analogue of the reference's
``df.rdd.mapPartitionsWithIndex(worker.train).collect()`` launch
(/root/reference/ddl_mnist_aztk.py:219 via dist-keras).  PySpark and a JVM are
not on this image, so that adapter is import-guarded and untested here.
"""
from __future__ import annotations

import multiprocessing as mp
import os
import pickle
import threading
import time
import traceback
from typing import Any, Callable, List, Optional

import cloudpickle


class BarrierStageError(RuntimeError):
    """A rank raised, died or timed out; carries the first remote traceback."""


class TaskContext:
    """What a task sees of the barrier stage."""

    def partition_id(self) -> int:  # pragma: no cover - interface
        raise NotImplementedError

    def num_tasks(self) -> int:  # pragma: no cover
        raise NotImplementedError

    def all_gather(self, message: str) -> List[str]:  # pragma: no cover
        raise NotImplementedError

    def barrier(self) -> None:  # pragma: no cover
        raise NotImplementedError

    def task_addresses(self) -> List[str]:  # pragma: no cover
        raise NotImplementedError


# ----------------------------------------------------------------------------
# local backend
# ----------------------------------------------------------------------------
class _LocalTaskContext(TaskContext):
    def __init__(self, pid: int, n: int, conn, timeout: float):
        self._pid, self._n, self._conn, self._timeout = pid, n, conn, timeout

    def partition_id(self):
        return self._pid

    def num_tasks(self):
        return self._n

    def all_gather(self, message: str = '') -> List[str]:
        self._conn.send(('gather', message))
        if not self._conn.poll(self._timeout):
            raise BarrierStageError(f'rank {self._pid}: barrier timed out after {self._timeout}s '
                                    f'(a peer is missing or dead)')
        kind, payload = self._conn.recv()
        if kind == 'abort':
            raise BarrierStageError(f'rank {self._pid}: stage aborted: {payload}')
        return payload

    def barrier(self):
        self.all_gather('')

    def task_addresses(self):
        return ['127.0.0.1'] * self._n


def _local_entry(pid: int, n: int, conn, task_blob: bytes, timeout: float, env: dict):
    os.environ.update(env)
    try:
        task = cloudpickle.loads(task_blob)
        ctx = _LocalTaskContext(pid, n, conn, timeout)
        result = task(ctx)
        conn.send(('result', cloudpickle.dumps(result)))
    except BaseException as e:  # noqa: BLE001 - everything must reach the driver
        try:
            conn.send(('error', f'{type(e).__name__}: {e}\n{traceback.format_exc()}'))
        except Exception:
            pass
    finally:
        try:
            conn.close()
        except Exception:
            pass


class LocalBarrierBackend:
    """All-or-nothing stage of ``n`` spawned processes on this host."""

    def __init__(self, default_parallelism: Optional[int] = None, timeout: float = 600.0,
                 start_method: str = 'spawn', stage_timeout: Optional[float] = None):
        """``timeout``: longest wait inside one barrier / all_gather for a missing peer.  ``stage_timeout``: optional
        bound on the whole stage (default: none — a training run may take hours)."""
        self._default = default_parallelism
        self.timeout = float(timeout)
        self.stage_timeout = float(stage_timeout) if stage_timeout else None
        self.start_method = start_method

    def default_parallelism(self) -> int:
        if self._default is not None:
            return self._default
        try:
            import torch
            if torch.cuda.is_available():
                return torch.cuda.device_count()
        except Exception:
            pass
        return 1

    def run_barrier_stage(self, n: int, task: Callable[[TaskContext], Any]) -> List[Any]:
        ctx = mp.get_context(self.start_method)
        blob = cloudpickle.dumps(task)
        pipes = [ctx.Pipe(duplex=True) for _ in range(n)]
        env = {'DDLS_LAUNCHER': 'local_barrier'}
        procs = [ctx.Process(target=_local_entry, args=(i, n, pipes[i][1], blob, self.timeout, env), daemon=False)
                 for i in range(n)]
        for p in procs:
            p.start()
        for _, child in pipes:
            child.close()
        conns = [p[0] for p in pipes]
        results: List[Any] = [None] * n
        done = [False] * n
        gather_buf: dict = {}
        error: Optional[str] = None
        # `timeout` bounds how long ranks may WAIT FOR EACH OTHER in a gather (a missing peer), not how long the stage
        # runs: a healthy long training run has no pipe traffic for hours.  Total runtime is bounded only by the optional
        # stage_timeout; rank death is caught by the liveness check below.
        gather_deadline: Optional[float] = None
        stage_deadline = (time.time() + self.stage_timeout) if self.stage_timeout else None
        try:
            while not all(done) and error is None:
                progressed = False
                for i, c in enumerate(conns):
                    if done[i]:
                        continue
                    try:
                        has = c.poll(0)
                    except (OSError, EOFError):
                        has = False
                    if has:
                        try:
                            kind, payload = c.recv()
                        except (EOFError, OSError):
                            error = f'rank {i} closed its channel without a result (exit code {procs[i].exitcode})'
                            break
                        progressed = True
                        if kind == 'gather':
                            if not gather_buf:
                                gather_deadline = time.time() + self.timeout
                            gather_buf[i] = payload
                            if len(gather_buf) == n - sum(done):
                                if sum(done):
                                    error = 'barrier entered after some ranks already returned'
                                    break
                                msgs = [gather_buf[j] for j in range(n)]
                                gather_buf.clear()
                                for cc in conns:
                                    cc.send(('gathered', msgs))
                                gather_deadline = None
                        elif kind == 'result':
                            results[i] = cloudpickle.loads(payload)
                            done[i] = True
                        elif kind == 'error':
                            error = f'rank {i} failed:\n{payload}'
                            break
                    elif not procs[i].is_alive():
                        # drain once more: the result may have landed between poll and is_alive
                        try:
                            if c.poll(0.05):
                                continue
                        except (OSError, EOFError):
                            pass
                        error = f'rank {i} died (exit code {procs[i].exitcode}) before returning'
                        break
                if error is None and not progressed:
                    now = time.time()
                    if gather_buf and gather_deadline is not None and now > gather_deadline:
                        error = f'barrier timed out after {self.timeout}s; ranks that never arrived: ' \
                                f'{[i for i in range(n) if not done[i] and i not in gather_buf]}'
                    elif stage_deadline is not None and now > stage_deadline:
                        error = f'stage timed out after {self.stage_timeout}s; unfinished ranks: ' \
                                f'{[i for i in range(n) if not done[i]]}'
                    else:
                        time.sleep(0.002)
        finally:
            if error is not None:
                for cc in conns:
                    try:
                        cc.send(('abort', error.splitlines()[0]))
                    except Exception:
                        pass
            for p in procs:
                p.join(timeout=0.5 if error else 30)
            for p in procs:
                if p.is_alive():
                    p.terminate()
                    p.join(timeout=5)
                if p.is_alive():
                    p.kill()
            for cc in conns:
                try:
                    cc.close()
                except Exception:
                    pass
        if error is not None:
            raise BarrierStageError(error)
        return results


class InProcessBackend:
    """Degenerate single-task backend (``num_proc=1`` without spawning): the reference's
    "SingleWorker" comparison (/root/reference/ddl_mnist_aztk.py:215)."""

    def default_parallelism(self) -> int:
        return 1

    def run_barrier_stage(self, n: int, task):
        if n != 1:
            raise ValueError('InProcessBackend runs exactly one task')

        class _Ctx(TaskContext):
            def partition_id(self): return 0
            def num_tasks(self): return 1
            def all_gather(self, message=''): return [message]
            def barrier(self): return None
            def task_addresses(self): return ['127.0.0.1']
        return [task(_Ctx())]


# ----------------------------------------------------------------------------
# real PySpark adapter (import-guarded: no pyspark / JVM on this image; tests/test_launcher.py drives it against a fake
# ``pyspark`` module that implements the six calls used here with threads)
# ----------------------------------------------------------------------------
class PySparkBarrierBackend:
    def __init__(self, spark_context=None):
        try:
            import pyspark  # noqa: F401
        except ImportError as e:
            raise ImportError('PySparkBarrierBackend needs pyspark and a JVM; neither is on this image. '
                              'Use LocalBarrierBackend.') from e
        from pyspark import SparkContext
        self.sc = spark_context or SparkContext.getOrCreate()

    def default_parallelism(self) -> int:
        return int(self.sc.defaultParallelism)

    def run_barrier_stage(self, n: int, task):
        blob = cloudpickle.dumps(task)

        def _mapper(_it):
            from pyspark import BarrierTaskContext
            bctx = BarrierTaskContext.get()

            class _Ctx(TaskContext):
                def partition_id(self): return bctx.partitionId()
                def num_tasks(self): return len(bctx.getTaskInfos())
                def all_gather(self, message=''): return list(bctx.allGather(message))
                def barrier(self): bctx.barrier()
                def task_addresses(self): return [t.address.split(':')[0] for t in bctx.getTaskInfos()]
            ctx = _Ctx()
            yield (ctx.partition_id(), cloudpickle.dumps(cloudpickle.loads(blob)(ctx)))

        pairs = self.sc.parallelize(range(n), n).barrier().mapPartitions(_mapper).collect()
        pairs.sort(key=lambda kv: kv[0])
        return [pickle.loads(v) for _, v in pairs]


def pyspark_available() -> bool:
    try:
        import pyspark  # noqa: F401
        return True
    except ImportError:
        return False
