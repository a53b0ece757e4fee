from .launcher import run
from .backends import (BarrierStageError, TaskContext, LocalBarrierBackend, InProcessBackend,
                       PySparkBarrierBackend, pyspark_available)

__all__ = ['run', 'BarrierStageError', 'TaskContext', 'LocalBarrierBackend', 'InProcessBackend',
           'PySparkBarrierBackend', 'pyspark_available']
