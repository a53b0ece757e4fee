"""NumPy float64 restatement of the Keras 2.1/2.2 math the reference scripts run.

This module is the numerical oracle for every CUDA kernel in the package and
the timed CPU baseline.  It restates — in plain NumPy, forward *and*
hand-derived backward — exactly the layers/optimizers/losses the reference
builds:

* MNIST CNN: Conv2D(32,3x3,valid)+ReLU, Conv2D(32,3x3)+ReLU, MaxPool 2x2,
  Flatten, Dense(225)+ReLU, Dense(10)+softmax, categorical cross-entropy
  (/root/reference/ddl_mnist_aztk.py:180-196).
* NYISO forecasters: GRU(128) / LSTM(128) over (25,1) windows + Dense(1),
  mean squared error (/root/reference/ddl_nyiso_aztk.py:201-207,249-255).
* Optimizers named by the reference: 'adam' (ddl_mnist_aztk.py:195,
  ddl_nyiso_aztk.py:255) and 'adagrad' (ddl_nyiso_aztk.py:207), in their
  Keras-2 forms.

Layouts are Keras layouts: Conv2D kernel (kh,kw,cin,cout), NHWC activations,
Dense kernel (in,out), LSTM gate order i,f,c,o, GRU gate order z,r,h with the
reset gate applied before the recurrent matmul, hard_sigmoid recurrent
activation.
"""
from __future__ import annotations

import numpy as np

EPS = 1e-7  # K.epsilon()


# ----------------------------------------------------------------------------
# activations
# ----------------------------------------------------------------------------
def hard_sigmoid(x):
    return np.clip(0.2 * x + 0.5, 0.0, 1.0)


def hard_sigmoid_grad(x):
    return np.where((x > -2.5) & (x < 2.5), 0.2, 0.0)


def relu(x):
    return np.maximum(x, 0.0)


def softmax(z):
    z = z - z.max(axis=-1, keepdims=True)
    e = np.exp(z)
    return e / e.sum(axis=-1, keepdims=True)


# ----------------------------------------------------------------------------
# losses
# ----------------------------------------------------------------------------
def categorical_crossentropy(probs, onehot):
    """Keras form on softmax outputs: renormalise, clip to [eps, 1-eps]."""
    p = probs / probs.sum(axis=-1, keepdims=True)
    p = np.clip(p, EPS, 1.0 - EPS)
    return float(np.mean(-np.sum(onehot * np.log(p), axis=-1)))


def mean_squared_error(pred, target):
    return float(np.mean(np.mean((pred - target) ** 2, axis=-1)))


# ----------------------------------------------------------------------------
# conv / pool / dense
# ----------------------------------------------------------------------------
def im2col_valid(x, kh, kw):
    """x: (B,H,W,C) -> (B,OH,OW,kh*kw*C), tap order (kh,kw,c)."""
    B, H, W, C = x.shape
    OH, OW = H - kh + 1, W - kw + 1
    cols = np.empty((B, OH, OW, kh, kw, C), dtype=x.dtype)
    for i in range(kh):
        for j in range(kw):
            cols[:, :, :, i, j, :] = x[:, i:i + OH, j:j + OW, :]
    return cols.reshape(B, OH, OW, kh * kw * C)


def conv2d_valid(x, w, b):
    kh, kw, cin, cout = w.shape
    cols = im2col_valid(x, kh, kw)
    return cols @ w.reshape(kh * kw * cin, cout) + b


def conv2d_valid_backward(x, w, dy):
    """Returns (dx, dw, db) for y = conv2d_valid(x, w, b)."""
    kh, kw, cin, cout = w.shape
    B, H, W, C = x.shape
    OH, OW = H - kh + 1, W - kw + 1
    cols = im2col_valid(x, kh, kw).reshape(-1, kh * kw * cin)
    dy2 = dy.reshape(-1, cout)
    dw = (cols.T @ dy2).reshape(w.shape)
    db = dy2.sum(axis=0)
    dcols = (dy2 @ w.reshape(kh * kw * cin, cout).T).reshape(B, OH, OW, kh, kw, cin)
    dx = np.zeros_like(x)
    for i in range(kh):
        for j in range(kw):
            dx[:, i:i + OH, j:j + OW, :] += dcols[:, :, :, i, j, :]
    return dx, dw, db


def maxpool2x2(x):
    """x: (B,H,W,C) with even H,W -> pooled (B,H/2,W/2,C) and arg index 0..3
    (first maximum wins, window order (dh,dw) row-major)."""
    B, H, W, C = x.shape
    xr = x.reshape(B, H // 2, 2, W // 2, 2, C).transpose(0, 1, 3, 2, 4, 5).reshape(B, H // 2, W // 2, 4, C)
    idx = xr.argmax(axis=3)
    out = np.take_along_axis(xr, idx[:, :, :, None, :], axis=3)[:, :, :, 0, :]
    return out, idx


def maxpool2x2_backward(dout, idx):
    B, PH, PW, C = dout.shape
    d = np.zeros((B, PH, PW, 4, C), dtype=dout.dtype)
    np.put_along_axis(d, idx[:, :, :, None, :], dout[:, :, :, None, :], axis=3)
    return d.reshape(B, PH, PW, 2, 2, C).transpose(0, 1, 3, 2, 4, 5).reshape(B, PH * 2, PW * 2, C)


# ----------------------------------------------------------------------------
# MNIST CNN (weights in Keras get_weights() order)
# ----------------------------------------------------------------------------
def mnist_cnn_forward(weights, x):
    """weights = [k1(3,3,1,32), b1, k2(3,3,32,32), b2, kd1(4608,225), bd1, kd2(225,10), bd2].
    x: (B,28,28,1).  Returns (probs, cache)."""
    k1, b1, k2, b2, kd1, bd1, kd2, bd2 = weights
    z1 = conv2d_valid(x, k1, b1)
    a1 = relu(z1)
    z2 = conv2d_valid(a1, k2, b2)
    a2 = relu(z2)
    p, idx = maxpool2x2(a2)
    flat = p.reshape(p.shape[0], -1)  # (h,w,c) order
    z3 = flat @ kd1 + bd1
    a3 = relu(z3)
    z4 = a3 @ kd2 + bd2
    probs = softmax(z4)
    cache = dict(x=x, z1=z1, a1=a1, z2=z2, a2=a2, idx=idx, pshape=p.shape, flat=flat, z3=z3, a3=a3, probs=probs)
    return probs, cache


def mnist_cnn_loss_and_grads(weights, x, onehot):
    """Mean categorical cross-entropy and gradients in get_weights() order.

    Uses the softmax+xent shortcut dz = (p - y)/B, which equals Keras' clipped
    form whenever no probability is clipped (always true away from 1e-7)."""
    k1, b1, k2, b2, kd1, bd1, kd2, bd2 = weights
    probs, c = mnist_cnn_forward(weights, x)
    B = x.shape[0]
    loss = categorical_crossentropy(probs, onehot)
    dz4 = (probs - onehot) / B
    gkd2 = c['a3'].T @ dz4
    gbd2 = dz4.sum(0)
    da3 = dz4 @ kd2.T
    dz3 = da3 * (c['z3'] > 0)
    gkd1 = c['flat'].T @ dz3
    gbd1 = dz3.sum(0)
    dflat = dz3 @ kd1.T
    dp = dflat.reshape(c['pshape'])
    da2 = maxpool2x2_backward(dp, c['idx'])
    dz2 = da2 * (c['z2'] > 0)
    da1, gk2, gb2 = conv2d_valid_backward(c['a1'], k2, dz2)
    dz1 = da1 * (c['z1'] > 0)
    _, gk1, gb1 = conv2d_valid_backward(c['x'], k1, dz1)
    return loss, [gk1, gb1, gk2, gb2, gkd1, gbd1, gkd2, gbd2], probs


# ----------------------------------------------------------------------------
# recurrent layers
# ----------------------------------------------------------------------------
def lstm_forward(kernel, recurrent, bias, x, h0=None, c0=None, return_sequences=False):
    """Keras LSTM, gate order i,f,c,o; hard_sigmoid gates, tanh cell.
    x: (B,T,D).  Returns (out, cache)."""
    B, T, D = x.shape
    H = recurrent.shape[0]
    h = np.zeros((B, H)) if h0 is None else h0
    c = np.zeros((B, H)) if c0 is None else c0
    hs, cs, zs = [h], [c], []
    for t in range(T):
        z = x[:, t, :] @ kernel + h @ recurrent + bias
        zi, zf, zc, zo = z[:, :H], z[:, H:2 * H], z[:, 2 * H:3 * H], z[:, 3 * H:]
        i, f, g, o = hard_sigmoid(zi), hard_sigmoid(zf), np.tanh(zc), hard_sigmoid(zo)
        c = f * c + i * g
        h = o * np.tanh(c)
        zs.append(z); hs.append(h); cs.append(c)
    out = np.stack(hs[1:], axis=1) if return_sequences else h
    return out, dict(x=x, hs=hs, cs=cs, zs=zs, H=H, return_sequences=return_sequences)


def lstm_backward(kernel, recurrent, bias, cache, dout):
    """dout: (B,H) for last-step output or (B,T,H) for sequences.
    Returns (dx, dkernel, drecurrent, dbias)."""
    x, hs, cs, zs, H = cache['x'], cache['hs'], cache['cs'], cache['zs'], cache['H']
    B, T, D = x.shape
    dk = np.zeros_like(kernel); dr = np.zeros_like(recurrent); db = np.zeros_like(bias)
    dx = np.zeros_like(x)
    dh = np.zeros((B, H)); dc = np.zeros((B, H))
    for t in reversed(range(T)):
        if cache['return_sequences']:
            dh = dh + dout[:, t, :]
        elif t == T - 1:
            dh = dh + dout
        z = zs[t]
        zi, zf, zc, zo = z[:, :H], z[:, H:2 * H], z[:, 2 * H:3 * H], z[:, 3 * H:]
        i, f, g, o = hard_sigmoid(zi), hard_sigmoid(zf), np.tanh(zc), hard_sigmoid(zo)
        c_prev, c_t = cs[t], cs[t + 1]
        tc = np.tanh(c_t)
        do = dh * tc
        dc = dc + dh * o * (1 - tc * tc)
        di = dc * g
        dg = dc * i
        df = dc * c_prev
        dz = np.concatenate([di * hard_sigmoid_grad(zi), df * hard_sigmoid_grad(zf),
                             dg * (1 - g * g), do * hard_sigmoid_grad(zo)], axis=1)
        dk += x[:, t, :].T @ dz
        dr += hs[t].T @ dz
        db += dz.sum(0)
        dx[:, t, :] = dz @ kernel.T
        dh = dz @ recurrent.T
        dc = dc * f
    return dx, dk, dr, db


def gru_forward(kernel, recurrent, bias, x, h0=None, return_sequences=False):
    """Keras GRU (reset_after=False), gate order z,r,h."""
    B, T, D = x.shape
    H = recurrent.shape[0]
    h = np.zeros((B, H)) if h0 is None else h0
    hs, saved = [h], []
    Uz, Ur, Uh = recurrent[:, :H], recurrent[:, H:2 * H], recurrent[:, 2 * H:]
    for t in range(T):
        xw = x[:, t, :] @ kernel + bias
        az = xw[:, :H] + h @ Uz
        ar = xw[:, H:2 * H] + h @ Ur
        z, r = hard_sigmoid(az), hard_sigmoid(ar)
        ah = xw[:, 2 * H:] + (r * h) @ Uh
        hh = np.tanh(ah)
        h = z * h + (1 - z) * hh
        saved.append((az, ar, ah))
        hs.append(h)
    out = np.stack(hs[1:], axis=1) if return_sequences else h
    return out, dict(x=x, hs=hs, saved=saved, H=H, return_sequences=return_sequences)


def gru_backward(kernel, recurrent, bias, cache, dout):
    x, hs, saved, H = cache['x'], cache['hs'], cache['saved'], cache['H']
    B, T, D = x.shape
    Uz, Ur, Uh = recurrent[:, :H], recurrent[:, H:2 * H], recurrent[:, 2 * H:]
    dk = np.zeros_like(kernel); dr_ = np.zeros_like(recurrent); db = np.zeros_like(bias)
    dx = np.zeros_like(x)
    dh = np.zeros((B, H))
    for t in reversed(range(T)):
        if cache['return_sequences']:
            dh = dh + dout[:, t, :]
        elif t == T - 1:
            dh = dh + dout
        az, ar, ah = saved[t]
        h_prev = hs[t]
        z, r, hh = hard_sigmoid(az), hard_sigmoid(ar), np.tanh(ah)
        dz = dh * (h_prev - hh)
        dhh = dh * (1 - z)
        dah = dhh * (1 - hh * hh)
        drh = dah @ Uh.T                      # grad wrt (r*h_prev)
        dr = drh * h_prev
        daz = dz * hard_sigmoid_grad(az)
        dar = dr * hard_sigmoid_grad(ar)
        da = np.concatenate([daz, dar, dah], axis=1)
        dk += x[:, t, :].T @ da
        db += da.sum(0)
        dr_[:, :H] += h_prev.T @ daz
        dr_[:, H:2 * H] += h_prev.T @ dar
        dr_[:, 2 * H:] += (r * h_prev).T @ dah
        dx[:, t, :] = da @ kernel.T
        dh = dh * z + drh * r + daz @ Uz.T + dar @ Ur.T
    return dx, dk, dr_, db


def rnn_regressor_loss_and_grads(cell, weights, x, y):
    """cell in {'lstm','gru'}; weights = [kernel, recurrent, bias, dense_k(H,1), dense_b(1)].
    Loss = Keras mean_squared_error averaged over the batch."""
    kernel, recurrent, bias, dk_, db_ = weights
    fwd, bwd = (lstm_forward, lstm_backward) if cell == 'lstm' else (gru_forward, gru_backward)
    h, cache = fwd(kernel, recurrent, bias, x)
    pred = h @ dk_ + db_
    loss = mean_squared_error(pred, y)
    B, O = pred.shape
    dpred = 2.0 * (pred - y) / (B * O)
    gdk = h.T @ dpred
    gdb = dpred.sum(0)
    dh = dpred @ dk_.T
    _, gk, gr, gb = bwd(kernel, recurrent, bias, cache, dh)
    return loss, [gk, gr, gb, gdk, gdb], pred


# ----------------------------------------------------------------------------
# optimizers (Keras 2.x update rules)
# ----------------------------------------------------------------------------
class KerasAdam:
    def __init__(self, lr=1e-3, beta_1=0.9, beta_2=0.999, epsilon=EPS):
        self.lr, self.b1, self.b2, self.eps = lr, beta_1, beta_2, epsilon
        self.t = 0
        self.m = None
        self.v = None

    def step(self, params, grads):
        if self.m is None:
            self.m = [np.zeros_like(p) for p in params]
            self.v = [np.zeros_like(p) for p in params]
        self.t += 1
        lr_t = self.lr * np.sqrt(1.0 - self.b2 ** self.t) / (1.0 - self.b1 ** self.t)
        for p, g, m, v in zip(params, grads, self.m, self.v):
            m[...] = self.b1 * m + (1 - self.b1) * g
            v[...] = self.b2 * v + (1 - self.b2) * g * g
            p -= lr_t * m / (np.sqrt(v) + self.eps)


class KerasAdagrad:
    def __init__(self, lr=0.01, epsilon=EPS):
        self.lr, self.eps = lr, epsilon
        self.a = None

    def step(self, params, grads):
        if self.a is None:
            self.a = [np.zeros_like(p) for p in params]
        for p, g, a in zip(params, grads, self.a):
            a += g * g
            p -= self.lr * g / (np.sqrt(a) + self.eps)


class KerasSGD:
    def __init__(self, lr=0.01):
        self.lr = lr

    def step(self, params, grads):
        for p, g in zip(params, grads):
            p -= self.lr * g
