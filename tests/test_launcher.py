"""spark.run contract tests against the local barrier backend (the real-PySpark adapter needs a JVM)."""
import os
import time

import pytest

from distributeddeeplearningspark_b200 import spark
from distributeddeeplearningspark_b200.spark import (BarrierStageError, InProcessBackend, LocalBarrierBackend,
                                                     pyspark_available)


def _ident(tag):
    return {'tag': tag, 'rank': int(os.environ['RANK']), 'size': int(os.environ['WORLD_SIZE']),
            'local_rank': int(os.environ['LOCAL_RANK']), 'master': os.environ['MASTER_ADDR'], 'pid': os.getpid()}


def test_run_returns_results_ordered_by_rank():
    out = spark.run(_ident, args=('t',), num_proc=3, backend=LocalBarrierBackend(timeout=120))
    assert [o['rank'] for o in out] == [0, 1, 2]
    assert all(o['size'] == 3 and o['tag'] == 't' and o['master'] == '127.0.0.1' for o in out)
    assert [o['local_rank'] for o in out] == [0, 1, 2]
    assert len({o['pid'] for o in out}) == 3


def test_inprocess_backend_single_rank():
    out = spark.run(lambda a, b=0: a + b, args=(2,), kwargs={'b': 5}, num_proc=1, backend=InProcessBackend())
    assert out == [7]


def _boom():
    if int(os.environ['RANK']) == 1:
        raise ValueError('rank one exploded')
    time.sleep(30)
    return 'late'


def test_first_exception_propagates_and_peers_are_torn_down():
    t0 = time.time()
    with pytest.raises(BarrierStageError, match='rank one exploded'):
        spark.run(_boom, num_proc=2, backend=LocalBarrierBackend(timeout=120))
    assert time.time() - t0 < 25          # did not wait for the sleeping peer


def _die():
    if int(os.environ['RANK']) == 0:
        os._exit(3)
    time.sleep(30)


def test_killed_rank_is_detected_not_hung():
    with pytest.raises(BarrierStageError, match='died|closed'):
        spark.run(_die, num_proc=2, backend=LocalBarrierBackend(timeout=120))


def _slow(seconds):
    time.sleep(seconds)
    return int(os.environ['RANK'])


def test_barrier_timeout_is_not_a_runtime_limit():
    """A task that runs longer than the barrier timeout without any pipe traffic is healthy, not hung."""
    out = spark.run(_slow, args=(4.0,), num_proc=2, backend=LocalBarrierBackend(timeout=1.5))
    assert out == [0, 1]


def test_stage_timeout_bounds_total_runtime_when_asked():
    with pytest.raises(BarrierStageError, match='stage timed out'):
        spark.run(_slow, args=(30.0,), num_proc=2, backend=LocalBarrierBackend(timeout=60, stage_timeout=3.0))


def _barrier_user():
    # tasks can keep using the stage barrier through torch.distributed after init
    import distributeddeeplearningspark_b200 as ddls
    ddls.init()
    ddls.barrier()
    return ddls.rank(), ddls.size()


def test_rank_size_through_init():
    out = spark.run(_barrier_user, num_proc=2, backend=LocalBarrierBackend(timeout=180), env={'DDLS_FORCE_CPU': '1'})
    assert out == [(0, 2), (1, 2)]


@pytest.mark.skipif(not pyspark_available(), reason='pyspark/JVM not on image')
def test_pyspark_adapter_contract():   # pragma: no cover
    from distributeddeeplearningspark_b200.spark import PySparkBarrierBackend
    out = spark.run(_ident, args=('s',), num_proc=2, backend=PySparkBarrierBackend())
    assert [o['rank'] for o in out] == [0, 1]


# ---- the PySpark adapter against a stand-in for the six PySpark calls it makes -------------------------------------
def _install_fake_pyspark(monkeypatch):
    """A minimal in-process ``pyspark``: SparkContext.getOrCreate / defaultParallelism / parallelize(...).barrier()
    .mapPartitions(f).collect(), and BarrierTaskContext.get() with partitionId / allGather / barrier / getTaskInfos.
    Barrier tasks run as threads (a real barrier stage runs them at the same time too)."""
    import sys
    import threading
    import types
    mod = types.ModuleType('pyspark')
    local = threading.local()

    class TaskInfo:
        def __init__(self, address): self.address = address

    class BarrierTaskContext:
        def __init__(self, pid, n, shared):
            self._pid, self._n, self._s = pid, n, shared

        @classmethod
        def get(cls): return local.ctx

        def partitionId(self): return self._pid

        def getTaskInfos(self): return [TaskInfo(f'127.0.0.1:{40000 + i}') for i in range(self._n)]

        def barrier(self): self._s['bar'].wait(timeout=30)

        def allGather(self, message=''):
            self._s['msgs'][self._pid] = message
            self._s['bar'].wait(timeout=30)
            out = list(self._s['msgs'])
            self._s['bar'].wait(timeout=30)
            return out

    class _Barrier:
        def __init__(self, parts): self.parts = parts

        def mapPartitions(self, f):
            parts = self.parts

            class _Collected:
                def collect(self_inner):
                    n = len(parts)
                    shared = {'bar': threading.Barrier(n), 'msgs': [None] * n}
                    out, errs = [None] * n, []

                    def run(i):
                        local.ctx = BarrierTaskContext(i, n, shared)
                        try:
                            out[i] = list(f(iter(parts[i])))
                        except BaseException as e:  # noqa: BLE001
                            errs.append(e); shared['bar'].abort()
                    ts = [threading.Thread(target=run, args=(i,)) for i in range(n)]
                    [t.start() for t in ts]; [t.join() for t in ts]
                    if errs:
                        raise errs[0]
                    return [x for part in reversed(out) for x in part]      # deliberately NOT in partition order
            return _Collected()

    class _RDD:
        def __init__(self, parts): self.parts = parts

        def barrier(self): return _Barrier(self.parts)

    class SparkContext:
        defaultParallelism = 3
        _inst = None

        @classmethod
        def getOrCreate(cls):
            cls._inst = cls._inst or cls()
            return cls._inst

        def parallelize(self, data, n):
            data = list(data)
            return _RDD([data[i::n] for i in range(n)])

    mod.SparkContext, mod.BarrierTaskContext = SparkContext, BarrierTaskContext
    monkeypatch.setitem(sys.modules, 'pyspark', mod)
    return mod


def _contract_task(ctx):
    ctx.barrier()
    return {'pid': ctx.partition_id(), 'n': ctx.num_tasks(), 'gathered': ctx.all_gather(f'hello from {ctx.partition_id()}'),
            'hosts': ctx.task_addresses()}


def test_pyspark_adapter_maps_the_barrier_stage_protocol(monkeypatch):
    """PySparkBarrierBackend against the stand-in: one task per partition in a barrier stage, BarrierTaskContext mapped to
    the launcher's TaskContext protocol, results re-ordered by partition id (collect() makes no ordering promise)."""
    _install_fake_pyspark(monkeypatch)
    from distributeddeeplearningspark_b200.spark import PySparkBarrierBackend
    be = PySparkBarrierBackend()
    assert be.default_parallelism() == 3
    out = be.run_barrier_stage(4, _contract_task)
    assert [o['pid'] for o in out] == [0, 1, 2, 3] and all(o['n'] == 4 for o in out)
    assert all(o['gathered'] == [f'hello from {i}' for i in range(4)] for o in out)
    assert all(o['hosts'] == ['127.0.0.1'] * 4 for o in out)


def _boom_task(ctx):
    if ctx.partition_id() == 2:
        raise ValueError('task two exploded')
    ctx.barrier()
    return 'ok'


def test_pyspark_adapter_propagates_a_task_failure(monkeypatch):
    _install_fake_pyspark(monkeypatch)
    from distributeddeeplearningspark_b200.spark import PySparkBarrierBackend
    with pytest.raises(Exception):
        PySparkBarrierBackend().run_barrier_stage(3, _boom_task)
