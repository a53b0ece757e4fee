from .misc import (get_os_username, determine_host_address, free_port, env_int, env_flag, env_str,
                   serialize_keras_model, deserialize_keras_model, new_dataframe_row, shuffle, precache,
                   to_one_hot_encoded_dense, history_executors_average, history_executor)
from .logging import get_logger, MetricsWriter
from .timing import PhaseTimers, Timeline, ClockSampler, flush_l2

__all__ = ['get_os_username', 'determine_host_address', 'free_port', 'env_int', 'env_flag', 'env_str',
           'serialize_keras_model', 'deserialize_keras_model', 'new_dataframe_row', 'shuffle', 'precache',
           'to_one_hot_encoded_dense', 'history_executors_average', 'history_executor',
           'get_logger', 'MetricsWriter', 'PhaseTimers', 'Timeline', 'ClockSampler', 'flush_l2']
