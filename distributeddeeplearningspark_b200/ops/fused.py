"""Pattern dispatch: recognise the reference's three architectures inside a ``Sequential`` and run them
through the fused sm_100a kernels (recurrent layer + fused head; CNN trunk + fused head)."""
from __future__ import annotations

from typing import Optional

import torch

from ..models import layers as L


def _plan(model):
    plan = getattr(model, '_fused_plan', None)
    if plan is not None:
        return plan
    ls = list(model.layers_list)
    kinds = [l.keras_class for l in ls]
    plan = {'kind': None}
    if len(ls) == 2 and kinds[0] in ('LSTM', 'GRU') and kinds[1] == 'Dense' and ls[1].activation in (None, 'linear') \
            and ls[0].units == 128 and not ls[0].return_sequences and model.input_shape[-1] == 1 \
            and ls[1].units <= 4:
        plan = {'kind': 'rnn_regressor', 'rnn': ls[0], 'dense': ls[1]}
    elif kinds == ['Conv2D', 'Activation', 'Conv2D', 'Activation', 'MaxPooling2D', 'Flatten', 'Dense', 'Activation',
                   'Dense', 'Activation'] \
            and ls[1].activation == 'relu' and ls[3].activation == 'relu' and ls[7].activation == 'relu' \
            and ls[9].activation == 'softmax' and ls[8].units <= 32:
        plan = {'kind': 'cnn', 'conv1': ls[0], 'conv2': ls[2], 'dense1': ls[6], 'dense2': ls[8]}
    model._fused_plan = plan
    return plan


def _cnn_trunk(plan, x):
    from . import cnn
    if cnn.native_available(plan, x):
        return cnn.trunk(plan, x)
    return None


def try_forward(model, x: torch.Tensor) -> Optional[torch.Tensor]:
    plan = _plan(model)
    if plan['kind'] == 'rnn_regressor':
        from . import heads
        h = plan['rnn'](x)
        if h.is_cuda and not torch.is_grad_enabled():
            return heads.dense_linear(h, plan['dense'].kernel, plan['dense'].bias)
        return plan['dense'](h)
    if plan['kind'] == 'cnn' and not torch.is_grad_enabled():
        from . import heads
        a = _cnn_trunk(plan, x)
        if a is not None:
            return heads.dense_softmax(a, plan['dense2'].kernel, plan['dense2'].bias)
    return None


def try_loss(model, x: torch.Tensor, y: torch.Tensor, loss_name: str) -> Optional[torch.Tensor]:
    plan = _plan(model)
    from . import heads
    if plan['kind'] == 'rnn_regressor' and loss_name in ('mean_squared_error', 'mse'):
        h = plan['rnn'](x)
        loss, _ = heads.dense_mse(h, plan['dense'].kernel, plan['dense'].bias, y)
        return loss
    if plan['kind'] == 'cnn' and loss_name == 'categorical_crossentropy':
        a = _cnn_trunk(plan, x)
        if a is None:   # trunk through torch ops, head still fused
            ls = list(model.layers_list)
            a = x
            for l in ls[:8]:
                a = l(a)
        loss, _ = heads.dense_softmax_xent(a, plan['dense2'].kernel, plan['dense2'].bias, y)
        return loss
    return None
