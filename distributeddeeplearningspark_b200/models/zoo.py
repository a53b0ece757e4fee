"""The reference's three networks plus the scaled stacked LSTM of BASELINE.json config 4.

* ``mnist_cnn``  — /root/reference/ddl_mnist_aztk.py:180-192 (1,048,853 parameters)
* ``nyiso_gru``  — /root/reference/ddl_nyiso_aztk.py:201-203 (50,049 parameters, ddl_nyiso_hdi.ipynb:549-553)
* ``nyiso_lstm`` — /root/reference/ddl_nyiso_aztk.py:249-251 (66,689 parameters, ddl_nyiso_hdi.ipynb:772-776)
* ``stacked_lstm`` — scaled forecaster (hidden 1024, seq 168, 2 layers): 12,596,225 parameters
"""
from __future__ import annotations

import torch

from .layers import Activation, Conv2D, Dense, Flatten, GRU, LSTM, MaxPooling2D
from .sequential import Sequential


def _seeded(fn, seed):
    if seed is None:
        return fn()
    state = torch.random.get_rng_state()
    torch.manual_seed(seed)
    try:
        return fn()
    finally:
        torch.random.set_rng_state(state)


def mnist_cnn(nb_classes: int = 10, nb_filters: int = 32, seed=None) -> Sequential:
    def build():
        m = Sequential()
        m.add(Conv2D(nb_filters, (3, 3), padding='valid', input_shape=(28, 28, 1)))
        m.add(Activation('relu'))
        m.add(Conv2D(nb_filters, (3, 3)))
        m.add(Activation('relu'))
        m.add(MaxPooling2D(pool_size=(2, 2)))
        m.add(Flatten())
        m.add(Dense(225))
        m.add(Activation('relu'))
        m.add(Dense(nb_classes))
        m.add(Activation('softmax'))
        m.compile(optimizer='adam', loss='categorical_crossentropy')
        return m
    return _seeded(build, seed)


def nyiso_gru(units: int = 128, len_seq: int = 25, len_out: int = 1, seed=None) -> Sequential:
    def build():
        m = Sequential()
        m.add(GRU(units, input_shape=(len_seq, 1)))
        m.add(Dense(len_out, activation='linear'))
        m.compile(optimizer='adagrad', loss='mean_squared_error')
        return m
    return _seeded(build, seed)


def nyiso_lstm(units: int = 128, len_seq: int = 25, len_out: int = 1, seed=None) -> Sequential:
    def build():
        m = Sequential()
        m.add(LSTM(units, input_shape=(len_seq, 1)))
        m.add(Dense(len_out, activation='linear'))
        m.compile(optimizer='adam', loss='mean_squared_error')
        return m
    return _seeded(build, seed)


def stacked_lstm(units: int = 1024, len_seq: int = 168, num_layers: int = 2, len_out: int = 1, seed=None) -> Sequential:
    def build():
        m = Sequential()
        for i in range(num_layers):
            kw = {'input_shape': (len_seq, 1)} if i == 0 else {}
            m.add(LSTM(units, return_sequences=(i < num_layers - 1), **kw))
        m.add(Dense(len_out, activation='linear'))
        m.compile(optimizer='adam', loss='mean_squared_error')
        return m
    return _seeded(build, seed)
