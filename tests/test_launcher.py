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
