from .state import (init, shutdown, is_initialized, rank, size, local_rank, local_size, device, topology,
                    communicator, Topology)

__all__ = ['init', 'shutdown', 'is_initialized', 'rank', 'size', 'local_rank', 'local_size', 'device',
           'topology', 'communicator', 'Topology']
