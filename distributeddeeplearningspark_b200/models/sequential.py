"""``Sequential`` container with the Keras surface the reference scripts use.

Covers ``Sequential()``, ``.add(layer)``, ``.summary()``, ``get_weights()`` /
``set_weights()`` and the JSON+weights serialisation dist-keras ships to its
workers (/root/reference/ddl_mnist_aztk.py:180-199,
/root/reference/ddl_nyiso_aztk.py:201-204,249-252).  The container is a torch
``nn.Module``; on CUDA, recognised layer runs are dispatched to fused
hand-written kernels (see ``ops.fused``).
"""
from __future__ import annotations

import json
from collections import Counter
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn as nn

from . import layers as L

_SNAKE = {'Conv2D': 'conv2d', 'Activation': 'activation', 'MaxPooling2D': 'max_pooling2d', 'Flatten': 'flatten',
          'Dense': 'dense', 'LSTM': 'lstm', 'GRU': 'gru'}

_LAYER_CLASSES = {c.keras_class: c for c in
                  (L.Conv2D, L.Activation, L.MaxPooling2D, L.Flatten, L.Dense, L.LSTM, L.GRU)}


class Sequential(nn.Module):
    def __init__(self, layers: Optional[Sequence[L.KerasLayer]] = None, name: str = 'sequential_1'):
        super().__init__()
        self.model_name = name
        self.layers_list = nn.ModuleList()
        self._shapes: List[Tuple[int, ...]] = []
        self._input_shape: Optional[Tuple[int, ...]] = None
        self._name_counter: Counter = Counter()
        self._fused_plan = None
        self.loss_name: Optional[str] = None
        self.optimizer_name: Optional[str] = None
        for l in (layers or []):
            self.add(l)

    # ------------------------------------------------------------------ build
    def add(self, layer: L.KerasLayer) -> None:
        if self._input_shape is None:
            if layer.input_shape_arg is None:
                raise ValueError('the first layer needs input_shape=...')
            self._input_shape = tuple(layer.input_shape_arg)
            cur = self._input_shape
        else:
            cur = self._shapes[-1]
        if layer.layer_name is None:
            base = _SNAKE.get(layer.keras_class, layer.keras_class.lower())
            self._name_counter[base] += 1
            layer.layer_name = f'{base}_{self._name_counter[base]}'
        layer.build(cur)
        self.layers_list.append(layer)
        self._shapes.append(tuple(layer.compute_output_shape(cur)))
        self._fused_plan = None

    @property
    def layers(self) -> List[L.KerasLayer]:
        return list(self.layers_list)

    @property
    def input_shape(self) -> Tuple[int, ...]:
        return self._input_shape

    @property
    def output_shape(self) -> Tuple[int, ...]:
        return self._shapes[-1]

    # ---------------------------------------------------------------- forward
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda:
            from ..ops import fused
            out = fused.try_forward(self, x)
            if out is not None:
                return out
        for l in self.layers_list:
            x = l(x)
        return x

    def compile(self, optimizer: Optional[str] = None, loss: Optional[str] = None, metrics=None) -> None:
        """Records names only; the trainers build the actual optimizer/loss."""
        self.optimizer_name = optimizer
        self.loss_name = loss

    # ---------------------------------------------------------------- weights
    def named_keras_weights(self) -> List[Tuple[str, nn.Parameter]]:
        out = []
        for l in self.layers_list:
            for suffix, p in l.keras_weights():
                out.append((f'{l.layer_name}/{suffix}', p))
        return out

    def get_weights(self) -> List[np.ndarray]:
        return [p.detach().cpu().numpy().copy() for _, p in self.named_keras_weights()]

    def set_weights(self, weights: Sequence[np.ndarray]) -> None:
        named = self.named_keras_weights()
        if len(named) != len(weights):
            raise ValueError(f'expected {len(named)} arrays, got {len(weights)}')
        with torch.no_grad():
            for (n, p), w in zip(named, weights):
                w = torch.as_tensor(np.asarray(w), dtype=p.dtype)
                if tuple(w.shape) != tuple(p.shape):
                    raise ValueError(f'{n}: shape {tuple(w.shape)} != {tuple(p.shape)}')
                p.copy_(w)

    def count_params(self) -> int:
        return sum(l.count_params() for l in self.layers_list)

    # ---------------------------------------------------------------- summary
    def summary(self, print_fn=print) -> str:
        width = 65
        lines = ['_' * width, f"{'Layer (type)':<29}{'Output Shape':<26}{'Param #':<10}", '=' * width]
        for i, (l, s) in enumerate(zip(self.layers_list, self._shapes)):
            shape = '(' + ', '.join(['None'] + [str(d) for d in s]) + ')'
            lines.append(f"{(l.layer_name + ' (' + l.keras_class + ')'):<29}{shape:<26}{l.count_params():<10}")
            lines.append(('=' if i == len(self.layers_list) - 1 else '_') * width)
        total = self.count_params()
        lines += [f'Total params: {total:,}', f'Trainable params: {total:,}', 'Non-trainable params: 0', '_' * width]
        text = '\n'.join(lines)
        if print_fn is not None:
            print_fn(text)
        return text

    # ---------------------------------------------------------- serialisation
    def get_config(self) -> dict:
        cfg = []
        for i, l in enumerate(self.layers_list):
            c = dict(l.get_config())
            c['name'] = l.layer_name
            if i == 0:
                c['input_shape'] = list(self._input_shape)
            cfg.append({'class_name': l.keras_class, 'config': c})
        return {'class_name': 'Sequential', 'name': self.model_name, 'config': cfg}

    def to_json(self) -> str:
        return json.dumps(self.get_config())

    @staticmethod
    def from_config(cfg: dict) -> 'Sequential':
        m = Sequential(name=cfg.get('name', 'sequential_1'))
        for item in cfg['config']:
            cls = _LAYER_CLASSES[item['class_name']]
            c = dict(item['config'])
            name = c.pop('name', None)
            ishape = c.pop('input_shape', None)
            c.pop('recurrent_activation', None)
            if cls in (L.LSTM, L.GRU):
                c.pop('activation', None)
            if cls in (L.Conv2D, L.Dense) and c.get('activation') == 'linear':
                c['activation'] = None
            if cls is L.Conv2D:
                c['kernel_size'] = tuple(c['kernel_size'])
            if cls is L.MaxPooling2D:
                c['pool_size'] = tuple(c['pool_size'])
            m.add(cls(**c, input_shape=ishape, name=name) if cls is not L.Flatten
                  else cls(input_shape=ishape, name=name))
        return m

    @staticmethod
    def from_json(text: str) -> 'Sequential':
        return Sequential.from_config(json.loads(text))


def serialize_model(model: Sequential) -> Dict:
    """dist-keras ``serialize_keras_model`` analogue: architecture JSON + weight list."""
    return {'model': model.to_json(), 'weights': model.get_weights(),
            'loss': model.loss_name, 'optimizer': model.optimizer_name}


def deserialize_model(blob: Dict) -> Sequential:
    m = Sequential.from_json(blob['model'])
    m.set_weights(blob['weights'])
    m.compile(optimizer=blob.get('optimizer'), loss=blob.get('loss'))
    return m


def clone_model(model: Sequential) -> Sequential:
    return deserialize_model(serialize_model(model))
