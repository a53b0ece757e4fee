"""Convergence evidence: the bf16-operand / fp32-accumulate / fp32-master kernels against the fp32 Keras math.

The reference trains in float32 on CPUs (/root/reference/ddl_mnist_aztk.py:195-196, ddl_nyiso_aztk.py:207, 255).  Every
tensor-core kernel here rounds its operands to bf16.  These tests train the same model from the same weights on the same
minibatches twice — through the hand-written kernels on the GPU and through the plain fp32 PyTorch restatement of the
Keras equations on the CPU — and require the two loss curves to track each other over 200 optimizer steps."""
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _curves(name, steps, batches, lr=None):
    import distributeddeeplearningspark_b200 as ddls
    from distributeddeeplearningspark_b200.trainers import StepEngine
    dev = torch.device('cuda', 0)
    out = {}
    for tag, device in (('kernels_bf16', dev), ('fp32_reference', torch.device('cpu'))):
        model = getattr(ddls.models, name)(seed=11)
        kw = {'optimizer_kwargs': {'lr': lr}} if lr else {}
        eng = StepEngine(model, optimizer=model.optimizer_name, loss=model.loss_name, device=device,
                         use_cuda_graph=(device.type == 'cuda'), **kw)
        losses = []
        for i in range(steps):
            x, y = batches[i % len(batches)]
            losses.append(float(eng.step(x.to(device), y.to(device))))
        out[tag] = losses
    os.makedirs('gpurun_out', exist_ok=True)
    with open(f'gpurun_out/convergence_{name}.json', 'w') as f:
        json.dump({'model': name, 'steps': steps, 'optimizer': model.optimizer_name, **out}, f)
    return np.asarray(out['kernels_bf16']), np.asarray(out['fp32_reference'])


def _smooth(a, w=10):
    return np.convolve(a, np.ones(w) / w, mode='valid')


def test_mnist_cnn_loss_curve_tracks_the_fp32_reference():
    from distributeddeeplearningspark_b200.data import mnist_like_arrays
    px, lab = mnist_like_arrays(128 * 25, seed=5)
    x = torch.from_numpy(px.astype(np.float32) / 250.0).reshape(-1, 28, 28, 1)
    y = torch.nn.functional.one_hot(torch.from_numpy(lab).long(), 10).float()
    batches = [(x[i:i + 128], y[i:i + 128]) for i in range(0, len(x), 128)]
    a, b = _curves('mnist_cnn', 200, batches)
    sa, sb = _smooth(a), _smooth(b)
    assert sb[-1] < 0.5 * sb[0], 'the fp32 reference itself did not learn'
    rel = np.abs(sa - sb) / np.maximum(sb, 0.05)
    assert rel.max() < 0.15 and rel.mean() < 0.05, (float(rel.max()), float(rel.mean()))


@pytest.mark.parametrize('name', ['nyiso_lstm', 'nyiso_gru'])
def test_nyiso_loss_curve_tracks_the_fp32_reference(name):
    from distributeddeeplearningspark_b200.data import nyiso_windows_arrays
    xs, ys = nyiso_windows_arrays(32 * 40, seed=3)
    x, y = torch.from_numpy(xs), torch.from_numpy(ys)
    batches = [(x[i:i + 32], y[i:i + 32]) for i in range(0, len(x), 32)]
    a, b = _curves(name, 200, batches)
    sa, sb = _smooth(a), _smooth(b)
    assert sb[-1] < 0.7 * sb[0], 'the fp32 reference itself did not learn'
    rel = np.abs(sa - sb) / np.maximum(sb, 1e-3)
    assert rel.max() < 0.25 and rel.mean() < 0.08, (float(rel.max()), float(rel.mean()))
