"""Keras-semantics layers as torch modules (parameters kept in Keras layouts).

The reference builds its networks from ``keras.layers`` (Conv2D, Activation,
MaxPooling2D, Flatten, Dense at /root/reference/ddl_mnist_aztk.py:180-192;
GRU/LSTM/Dense at /root/reference/ddl_nyiso_aztk.py:201-203,249-251).  These
modules keep the same constructor arguments, default initialisers and weight
layouts (``get_weights()`` order) so a trained model round-trips through the
Keras-layout checkpoint, but run on torch tensors.  On a CUDA device the
forward/backward of the hot layers is routed to the hand-written sm_100a
kernels in ``ops``; on CPU the same equations are expressed with torch
primitives (this is what the CPU tests and the gloo plumbing tests exercise).
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


def hard_sigmoid(x: torch.Tensor) -> torch.Tensor:
    """Keras-2 ``hard_sigmoid``: clip(0.2*x + 0.5, 0, 1)."""
    return torch.clamp(0.2 * x + 0.5, 0.0, 1.0)


def glorot_uniform_(t: torch.Tensor, fan_in: int, fan_out: int, generator=None) -> torch.Tensor:
    limit = math.sqrt(6.0 / (fan_in + fan_out))
    with torch.no_grad():
        t.uniform_(-limit, limit, generator=generator)
    return t


def orthogonal_(t: torch.Tensor, generator=None) -> torch.Tensor:
    """Keras ``Orthogonal`` initialiser for a (rows, cols) matrix."""
    rows, cols = t.shape
    a = torch.randn(max(rows, cols), min(rows, cols), generator=generator, dtype=torch.float64)
    q, r = torch.linalg.qr(a)
    q = q * torch.sign(torch.diagonal(r)).unsqueeze(0)
    if rows < cols:
        q = q.t()
    with torch.no_grad():
        t.copy_(q[:rows, :cols].to(t.dtype))
    return t


class KerasLayer(nn.Module):
    """Base: tracks Keras weight names/order and output-shape inference."""

    keras_class = 'Layer'

    def __init__(self, input_shape: Optional[Sequence[int]] = None, name: Optional[str] = None):
        super().__init__()
        self.input_shape_arg = tuple(input_shape) if input_shape is not None else None
        self.layer_name = name
        self.built = False

    # -- to be overridden -----------------------------------------------
    def build(self, input_shape: Tuple[int, ...]) -> None:
        self.built = True

    def compute_output_shape(self, input_shape: Tuple[int, ...]) -> Tuple[int, ...]:
        return tuple(input_shape)

    def keras_weights(self) -> List[Tuple[str, nn.Parameter]]:
        """[(suffix, parameter)] in Keras ``get_weights()`` order."""
        return []

    def get_config(self) -> dict:
        return {}

    def count_params(self) -> int:
        return sum(p.numel() for _, p in self.keras_weights())


class Activation(KerasLayer):
    keras_class = 'Activation'

    def __init__(self, activation: str, **kw):
        super().__init__(**kw)
        self.activation = activation

    def forward(self, x):
        return apply_activation(x, self.activation)

    def get_config(self):
        return {'activation': self.activation}


def apply_activation(x, name: Optional[str]):
    if name in (None, 'linear'):
        return x
    if name == 'relu':
        return F.relu(x)
    if name == 'softmax':
        return F.softmax(x, dim=-1)
    if name == 'tanh':
        return torch.tanh(x)
    if name == 'sigmoid':
        return torch.sigmoid(x)
    if name == 'hard_sigmoid':
        return hard_sigmoid(x)
    raise ValueError(f'unsupported activation {name!r}')


class Conv2D(KerasLayer):
    """``Conv2D(filters, (kh,kw), padding='valid')`` on NHWC tensors; kernel (kh,kw,cin,cout)."""

    keras_class = 'Conv2D'

    def __init__(self, filters: int, kernel_size, padding: str = 'valid', activation: Optional[str] = None,
                 input_shape=None, name=None):
        super().__init__(input_shape=input_shape, name=name)
        if isinstance(kernel_size, int):
            kernel_size = (kernel_size, kernel_size)
        if padding != 'valid':
            raise ValueError("only padding='valid' is used by the reference (ddl_mnist_aztk.py:181-183)")
        self.filters = int(filters)
        self.kernel_size = tuple(int(k) for k in kernel_size)
        self.padding = padding
        self.activation = activation
        self.kernel: Optional[nn.Parameter] = None
        self.bias: Optional[nn.Parameter] = None

    def build(self, input_shape):
        cin = int(input_shape[-1])
        kh, kw = self.kernel_size
        self.kernel = nn.Parameter(torch.empty(kh, kw, cin, self.filters))
        self.bias = nn.Parameter(torch.zeros(self.filters))
        glorot_uniform_(self.kernel, kh * kw * cin, kh * kw * self.filters)
        self.built = True

    def compute_output_shape(self, s):
        kh, kw = self.kernel_size
        return (s[0] - kh + 1, s[1] - kw + 1, self.filters)

    def keras_weights(self):
        return [('kernel:0', self.kernel), ('bias:0', self.bias)]

    def forward(self, x):
        w = self.kernel.permute(3, 2, 0, 1)
        y = F.conv2d(x.permute(0, 3, 1, 2), w, self.bias).permute(0, 2, 3, 1)
        return apply_activation(y, self.activation)

    def get_config(self):
        return {'filters': self.filters, 'kernel_size': list(self.kernel_size), 'padding': self.padding,
                'activation': self.activation or 'linear'}


class MaxPooling2D(KerasLayer):
    keras_class = 'MaxPooling2D'

    def __init__(self, pool_size=(2, 2), **kw):
        super().__init__(**kw)
        if isinstance(pool_size, int):
            pool_size = (pool_size, pool_size)
        self.pool_size = tuple(pool_size)

    def compute_output_shape(self, s):
        return (s[0] // self.pool_size[0], s[1] // self.pool_size[1], s[2])

    def forward(self, x):
        return F.max_pool2d(x.permute(0, 3, 1, 2), self.pool_size).permute(0, 2, 3, 1)

    def get_config(self):
        return {'pool_size': list(self.pool_size)}


class Flatten(KerasLayer):
    keras_class = 'Flatten'

    def compute_output_shape(self, s):
        return (int(np.prod(s)),)

    def forward(self, x):
        return x.reshape(x.shape[0], -1)  # NHWC => (h,w,c) order like Keras


class Dense(KerasLayer):
    keras_class = 'Dense'

    def __init__(self, units: int, activation: Optional[str] = None, input_shape=None, name=None):
        super().__init__(input_shape=input_shape, name=name)
        self.units = int(units)
        self.activation = activation
        self.kernel = None
        self.bias = None

    def build(self, input_shape):
        fin = int(input_shape[-1])
        self.kernel = nn.Parameter(torch.empty(fin, self.units))
        self.bias = nn.Parameter(torch.zeros(self.units))
        glorot_uniform_(self.kernel, fin, self.units)
        self.built = True

    def compute_output_shape(self, s):
        return tuple(s[:-1]) + (self.units,)

    def keras_weights(self):
        return [('kernel:0', self.kernel), ('bias:0', self.bias)]

    def forward(self, x):
        return apply_activation(x @ self.kernel + self.bias, self.activation)

    def get_config(self):
        return {'units': self.units, 'activation': self.activation or 'linear'}


class _Recurrent(KerasLayer):
    n_gates = 0

    def __init__(self, units: int, return_sequences: bool = False, input_shape=None, name=None):
        super().__init__(input_shape=input_shape, name=name)
        self.units = int(units)
        self.return_sequences = bool(return_sequences)
        self.kernel = None
        self.recurrent_kernel = None
        self.bias = None

    def build(self, input_shape):
        d = int(input_shape[-1])
        H, G = self.units, self.n_gates
        self.kernel = nn.Parameter(torch.empty(d, G * H))
        self.recurrent_kernel = nn.Parameter(torch.empty(H, G * H))
        self.bias = nn.Parameter(torch.zeros(G * H))
        glorot_uniform_(self.kernel, d, G * H)
        # Keras draws one orthogonal (H, G*H) matrix
        orthogonal_(self.recurrent_kernel)
        self._init_bias()
        self.built = True

    def _init_bias(self):
        pass

    def compute_output_shape(self, s):
        return (s[0], self.units) if self.return_sequences else (self.units,)

    def keras_weights(self):
        return [('kernel:0', self.kernel), ('recurrent_kernel:0', self.recurrent_kernel), ('bias:0', self.bias)]

    def get_config(self):
        return {'units': self.units, 'return_sequences': self.return_sequences,
                'activation': 'tanh', 'recurrent_activation': 'hard_sigmoid'}


class LSTM(_Recurrent):
    """Keras-2 LSTM: gates i,f,c,o; hard_sigmoid gates; unit forget bias."""

    keras_class = 'LSTM'
    n_gates = 4

    def _init_bias(self):
        with torch.no_grad():
            self.bias[self.units:2 * self.units] = 1.0

    def forward(self, x):
        from ..ops import recurrent as rops
        if rops.native_available(x, self.units):
            return rops.lstm_layer(x, self.kernel, self.recurrent_kernel, self.bias, self.return_sequences)
        return lstm_reference(x, self.kernel, self.recurrent_kernel, self.bias, self.return_sequences)


class GRU(_Recurrent):
    """Keras-2 GRU (reset_after=False): gates z,r,h; reset applied before the recurrent matmul."""

    keras_class = 'GRU'
    n_gates = 3

    def forward(self, x):
        from ..ops import recurrent as rops
        if rops.native_available(x, self.units):
            return rops.gru_layer(x, self.kernel, self.recurrent_kernel, self.bias, self.return_sequences)
        return gru_reference(x, self.kernel, self.recurrent_kernel, self.bias, self.return_sequences)


def lstm_reference(x, kernel, recurrent, bias, return_sequences=False):
    """Torch-primitive LSTM with the Keras equations (autograd-differentiable)."""
    B, T, _ = x.shape
    H = recurrent.shape[0]
    h = x.new_zeros(B, H)
    c = x.new_zeros(B, H)
    xw = x @ kernel + bias  # (B,T,4H)
    outs = []
    for t in range(T):
        z = xw[:, t] + h @ recurrent
        i = hard_sigmoid(z[:, :H]); f = hard_sigmoid(z[:, H:2 * H])
        g = torch.tanh(z[:, 2 * H:3 * H]); o = hard_sigmoid(z[:, 3 * H:])
        c = f * c + i * g
        h = o * torch.tanh(c)
        if return_sequences:
            outs.append(h)
    return torch.stack(outs, dim=1) if return_sequences else h


def gru_reference(x, kernel, recurrent, bias, return_sequences=False):
    B, T, _ = x.shape
    H = recurrent.shape[0]
    h = x.new_zeros(B, H)
    xw = x @ kernel + bias
    Uzr, Uh = recurrent[:, :2 * H], recurrent[:, 2 * H:]
    outs = []
    for t in range(T):
        azr = xw[:, t, :2 * H] + h @ Uzr
        z = hard_sigmoid(azr[:, :H]); r = hard_sigmoid(azr[:, H:])
        hh = torch.tanh(xw[:, t, 2 * H:] + (r * h) @ Uh)
        h = z * h + (1 - z) * hh
        if return_sequences:
            outs.append(h)
    return torch.stack(outs, dim=1) if return_sequences else h
