"""On-device MNIST preprocessing: MinMax (o_max = 250 like /root/reference/ddl_mnist_aztk.py:133-134),
reshape to (28,28,1) and one-hot labels in one pass over the uint8 pixels that crossed PCIe."""
from __future__ import annotations

import torch

from .. import _native as N


def mnist_preprocess(pixels_u8: torch.Tensor, labels_i64: torch.Tensor, classes: int = 10, o_min: float = 0.0,
                     o_max: float = 250.0, n_min: float = 0.0, n_max: float = 1.0, out_x=None, out_y=None):
    n, npix = pixels_u8.shape
    side = int(round(npix ** 0.5))
    if pixels_u8.is_cuda:
        x = out_x if out_x is not None else torch.empty((n, side, side, 1), dtype=torch.float32, device=pixels_u8.device)
        y = out_y if out_y is not None else torch.empty((n, classes), dtype=torch.float32, device=pixels_u8.device)
        N.call('ddls_mnist_preprocess', pixels_u8.data_ptr(), labels_i64.data_ptr(), x.data_ptr(), y.data_ptr(), n,
               npix, classes, o_min, (n_max - n_min) / (o_max - o_min), n_min, N.stream_ptr(pixels_u8.device))
        return x, y
    x = ((pixels_u8.float() - o_min) * ((n_max - n_min) / (o_max - o_min)) + n_min).reshape(n, side, side, 1)
    y = torch.nn.functional.one_hot(labels_i64, classes).float()
    return x, y
