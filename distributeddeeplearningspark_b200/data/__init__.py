from .frame import DataFrame, Row, VectorAssembler
from .transformers import (Transformer, MinMaxTransformer, StandardTransformer, OneHotTransformer,
                           BinaryLabelTransformer, ReshapeTransformer, DenseTransformer, LabelIndexTransformer)
from .window import Window, lag, lead, make_nyiso_windows, split_train_test
from .synthetic import mnist_like_arrays, mnist_like_frame, nyiso_like_frame, nyiso_windows_arrays
from .feeder import DeviceFeeder
from .arrow_feed import ArrowStreamFeeder, ipc_file_source, write_partition_file

__all__ = ['DataFrame', 'Row', 'VectorAssembler', 'Transformer', 'MinMaxTransformer', 'StandardTransformer',
           'OneHotTransformer', 'BinaryLabelTransformer', 'ReshapeTransformer', 'DenseTransformer',
           'LabelIndexTransformer', 'Window', 'lag', 'lead', 'make_nyiso_windows', 'split_train_test',
           'mnist_like_arrays', 'mnist_like_frame', 'nyiso_like_frame', 'nyiso_windows_arrays', 'DeviceFeeder',
           'ArrowStreamFeeder', 'ipc_file_source', 'write_partition_file']
