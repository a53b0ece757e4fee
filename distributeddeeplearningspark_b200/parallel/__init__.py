from .comm import ReduceOp, Average, Sum, Communicator, LocalComm, TorchDistComm, make_communicator
from .distributed_optimizer import DistributedOptimizer, Compression, plan_buckets
from .functions import (allreduce, allreduce_, allgather, broadcast, broadcast_, barrier, allgather_object,
                        broadcast_object, broadcast_parameters, broadcast_optimizer_state)

__all__ = ['ReduceOp', 'Average', 'Sum', 'Communicator', 'LocalComm', 'TorchDistComm', 'make_communicator',
           'DistributedOptimizer', 'Compression', 'plan_buckets', 'allreduce', 'allreduce_', 'allgather',
           'broadcast', 'broadcast_', 'barrier', 'allgather_object', 'broadcast_object',
           'broadcast_parameters', 'broadcast_optimizer_state']
