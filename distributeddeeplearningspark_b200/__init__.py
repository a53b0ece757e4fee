"""distributeddeeplearningspark_b200 — B200-native, Spark-launched data-parallel training.

Top-level surface (Horovod-shaped core + dist-keras-shaped trainers):

    import distributeddeeplearningspark_b200 as ddls
    ddls.spark.run(fn, num_proc=8)                     # barrier-stage launcher, one rank per GPU
    ddls.init(); ddls.rank(); ddls.size(); ddls.local_rank()
    ddls.broadcast_parameters(model, root_rank=0)
    opt = ddls.DistributedOptimizer(ddls.optim.KerasAdam(model.parameters()))
    ddls.trainers.ADAG(keras_model=..., ...).train(df)  # the reference scripts' API
"""
from . import runtime, spark, parallel, optim, models, data, trainers, ckpt, utils, oracle  # noqa: F401
from .runtime import init, shutdown, is_initialized, rank, size, local_rank, local_size, device, topology
from .parallel import (Average, Sum, Compression, DistributedOptimizer, allreduce, allreduce_, allgather,
                       broadcast, broadcast_, barrier, broadcast_parameters, broadcast_optimizer_state,
                       allgather_object, broadcast_object)

__version__ = '0.1.0'

__all__ = ['runtime', 'spark', 'parallel', 'optim', 'models', 'data', 'trainers', 'ckpt', 'utils', 'oracle',
           'init', 'shutdown', 'is_initialized', 'rank', 'size', 'local_rank', 'local_size', 'device', 'topology',
           'Average', 'Sum', 'Compression', 'DistributedOptimizer', 'allreduce', 'allreduce_', 'allgather',
           'broadcast', 'broadcast_', 'barrier', 'broadcast_parameters', 'broadcast_optimizer_state',
           'allgather_object', 'broadcast_object']
