from .layers import (Activation, Conv2D, Dense, Flatten, GRU, LSTM, MaxPooling2D, KerasLayer, hard_sigmoid,
                     lstm_reference, gru_reference)
from .sequential import Sequential, serialize_model, deserialize_model, clone_model
from .zoo import mnist_cnn, nyiso_gru, nyiso_lstm, stacked_lstm

__all__ = ['Activation', 'Conv2D', 'Dense', 'Flatten', 'GRU', 'LSTM', 'MaxPooling2D', 'KerasLayer', 'Sequential',
           'serialize_model', 'deserialize_model', 'clone_model', 'mnist_cnn', 'nyiso_gru', 'nyiso_lstm',
           'stacked_lstm', 'hard_sigmoid', 'lstm_reference', 'gru_reference']
