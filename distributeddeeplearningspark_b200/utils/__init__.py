from .misc import get_os_username, determine_host_address, free_port, env_int, env_flag, env_str
from .logging import get_logger, MetricsWriter
from .timing import CudaTimer, PhaseTimers, Timeline, ClockSampler, flush_l2

__all__ = ['get_os_username', 'determine_host_address', 'free_port', 'env_int', 'env_flag', 'env_str',
           'get_logger', 'MetricsWriter', 'CudaTimer', 'PhaseTimers', 'Timeline', 'ClockSampler', 'flush_l2']
