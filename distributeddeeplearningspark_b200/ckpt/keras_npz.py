"""Keras-layout checkpoints: ``.npz`` keyed by Keras weight names + a JSON architecture sidecar.

The reference never saves a model (the trained Keras object only lives in driver memory,
/root/reference/ddl_mnist_aztk.py:219); BASELINE.json asks for "rank-0 saves the trained Keras-style
CNN and LSTM weights".  The layout is therefore Keras' own ``get_weights()`` order and array shapes
(SURVEY.md §5): CNN ``[conv2d_1/kernel:0 (3,3,1,32), conv2d_1/bias:0, conv2d_2/kernel:0 (3,3,32,32), ...,
dense_1/kernel:0 (4608,225) rows in (h,w,c) order, ...]``; LSTM ``[lstm_1/kernel:0 (1,512),
lstm_1/recurrent_kernel:0 (128,512), lstm_1/bias:0 (512)]`` gate order i,f,c,o; GRU z,r,h.  HDF5 is
not available on the image (no h5py), hence ``.npz``.  Optimizer slots and the step count are stored
under ``optimizer/...`` keys so a resumed run continues bit-identically.
"""
from __future__ import annotations

import json
import os
from typing import Dict, Optional

import numpy as np
import torch

from ..models.sequential import Sequential

FORMAT_VERSION = 1


def _paths(path: str):
    base = path[:-4] if path.endswith('.npz') else path
    return base + '.npz', base + '.json'


def save_weights(path: str, model: Sequential) -> str:
    npz, js = _paths(path)
    arrays = {name: p.detach().cpu().numpy() for name, p in model.named_keras_weights()}
    tmp = npz + '.tmp.npz'
    np.savez(tmp, **arrays)
    os.replace(tmp, npz)
    with open(js, 'w') as f:
        json.dump({'format_version': FORMAT_VERSION, 'architecture': model.get_config(),
                   'weight_names': list(arrays.keys()), 'loss': model.loss_name,
                   'optimizer': model.optimizer_name}, f, indent=1)
    return npz


def load_weights(path: str, model: Optional[Sequential] = None) -> Sequential:
    npz, js = _paths(path)
    with open(js) as f:
        meta = json.load(f)
    if model is None:
        model = Sequential.from_config(meta['architecture'])
        model.compile(optimizer=meta.get('optimizer'), loss=meta.get('loss'))
    with np.load(npz) as z:
        model.set_weights([z[n] for n in meta['weight_names']])
    return model


def save_checkpoint(path: str, model: Sequential, optimizer=None, step: Optional[int] = None,
                    extra: Optional[Dict] = None, rank: Optional[int] = None) -> Optional[str]:
    """Rank 0 writes weights (+ optimizer slots, step count).  Other ranks return None."""
    if rank is None:
        from .. import runtime
        rank = runtime.rank() if runtime.is_initialized() else 0
    inner0 = getattr(optimizer, 'inner', optimizer)
    if inner0 is not None and hasattr(inner0, 'gather_slots'):
        inner0.gather_slots()           # collective: slots sharded by the data-parallel step kernel are exchanged first
    if rank != 0:
        return None
    npz, js = _paths(path)
    arrays = {name: p.detach().cpu().numpy() for name, p in model.named_keras_weights()}
    meta = {'format_version': FORMAT_VERSION, 'architecture': model.get_config(),
            'weight_names': list(arrays.keys()), 'loss': model.loss_name, 'optimizer': model.optimizer_name,
            'step': step, 'extra': extra or {}}
    inner = getattr(optimizer, 'inner', optimizer)
    if inner is not None and hasattr(inner, 'slots'):
        sd = inner.state_dict()
        meta['optimizer_state'] = {'kind': sd['kind'], 'hyper': sd['hyper'], 'step': sd['step'],
                                   'slots': list(sd['slots'].keys())}
        for k, v in sd['slots'].items():
            arrays[f'optimizer/{k}'] = v.numpy()
    tmp = npz + '.tmp.npz'
    np.savez(tmp, **arrays)
    os.replace(tmp, npz)
    with open(js, 'w') as f:
        json.dump(meta, f, indent=1)
    return npz


def load_checkpoint(path: str, model: Optional[Sequential] = None, optimizer=None):
    """Returns (model, meta).  Restores optimizer slots when ``optimizer`` is given."""
    npz, js = _paths(path)
    with open(js) as f:
        meta = json.load(f)
    if model is None:
        model = Sequential.from_config(meta['architecture'])
        model.compile(optimizer=meta.get('optimizer'), loss=meta.get('loss'))
    with np.load(npz) as z:
        model.set_weights([z[n] for n in meta['weight_names']])
        inner = getattr(optimizer, 'inner', optimizer)
        if inner is not None and 'optimizer_state' in meta:
            os_ = meta['optimizer_state']
            inner.load_state_dict({'kind': os_['kind'], 'hyper': os_['hyper'], 'step': os_['step'],
                                   'slots': {k: torch.from_numpy(z[f'optimizer/{k}']) for k in os_['slots']}})
    return model, meta


def resume(path: str, model: Sequential, optimizer=None, root_rank: int = 0) -> Dict:
    """Load on ``root_rank`` and broadcast parameters + optimizer state to every rank."""
    from .. import runtime
    from ..parallel import broadcast_object, broadcast_optimizer_state, broadcast_parameters
    meta = {}
    if not runtime.is_initialized() or runtime.rank() == root_rank:
        _, meta = load_checkpoint(path, model, optimizer)
    if runtime.is_initialized() and runtime.size() > 1:
        meta = broadcast_object(meta, root_rank)
        inner = getattr(optimizer, 'inner', optimizer)
        if inner is not None and hasattr(inner, 'flat'):
            broadcast_parameters(inner, root_rank)
            broadcast_optimizer_state(inner, root_rank)
        else:
            broadcast_parameters(model, root_rank)
            if optimizer is not None:
                broadcast_optimizer_state(optimizer, root_rank)
    return meta
