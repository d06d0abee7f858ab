"""Oracle: depthwise k x k convolution (and the 1x1 pointwise convolution) of EfficientNet's
depthwise-separable blocks, written out as explicit sums in numpy float64.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Parity: the arithmetic lives in PyTorch itself —
timm's `efficientnet_b2` (un-vendored; built at
/root/reference/previous/V3-efb2/train_advanced_v3.0.py:330) calls `F.conv2d(x, w, stride=s,
padding=k//2, groups=C)` for `conv_dw` and `F.conv2d(x, w)` with a 1x1 kernel for
`conv_pw` / `conv_pwl`.  PyTorch is installed here, so this restatement is PINNED against
`torch.nn.functional.conv2d` and its autograd in float64 on CPU (tests/test_host_logic.py) and
that is what the GPU tests hold the kernels to.
"""
from __future__ import annotations

import numpy as np


def out_size(n: int, k: int, stride: int) -> int:
    return (n + 2 * (k // 2) - k) // stride + 1


def dwconv_fwd_ref(x: np.ndarray, w: np.ndarray, stride: int) -> np.ndarray:
    """x [B,H,W,C], w [k,k,C] -> y [B,Ho,Wo,C]:
    y[b,oy,ox,c] = sum_{kh,kw} xpad[b, oy*s + kh, ox*s + kw, c] * w[kh,kw,c]."""
    x = np.asarray(x, np.float64)
    w = np.asarray(w, np.float64)
    B, H, W, C = x.shape
    k = w.shape[0]
    p = k // 2
    Ho, Wo = out_size(H, k, stride), out_size(W, k, stride)
    xp = np.zeros((B, H + 2 * p, W + 2 * p, C))
    xp[:, p:p + H, p:p + W] = x
    y = np.zeros((B, Ho, Wo, C))
    for kh in range(k):
        for kw in range(k):
            y += xp[:, kh:kh + (Ho - 1) * stride + 1:stride,
                    kw:kw + (Wo - 1) * stride + 1:stride] * w[kh, kw]
    return y


def dwconv_bwd_ref(dy: np.ndarray, x: np.ndarray, w: np.ndarray, stride: int):
    """-> (dx [B,H,W,C], dw [k,k,C]): the adjoints of dwconv_fwd_ref in x and in w."""
    dy = np.asarray(dy, np.float64)
    x = np.asarray(x, np.float64)
    w = np.asarray(w, np.float64)
    B, H, W, C = x.shape
    k = w.shape[0]
    p = k // 2
    Ho, Wo = dy.shape[1], dy.shape[2]
    xp = np.zeros((B, H + 2 * p, W + 2 * p, C))
    xp[:, p:p + H, p:p + W] = x
    dxp = np.zeros_like(xp)
    dw = np.zeros((k, k, C))
    for kh in range(k):
        for kw in range(k):
            sl = (slice(None), slice(kh, kh + (Ho - 1) * stride + 1, stride),
                  slice(kw, kw + (Wo - 1) * stride + 1, stride))
            dw[kh, kw] = (xp[sl] * dy).sum(axis=(0, 1, 2))
            dxp[sl] += dy * w[kh, kw]
    return dxp[:, p:p + H, p:p + W], dw


def pwconv_fwd_ref(x: np.ndarray, w: np.ndarray, b=None) -> np.ndarray:
    """x [B,H,W,Cin], w [Cout,Cin] -> [B,H,W,Cout] (a 1x1 convolution is a matrix product)."""
    y = np.asarray(x, np.float64) @ np.asarray(w, np.float64).T
    return y if b is None else y + np.asarray(b, np.float64)


# ----------------------------------------------------------------------- BatchNorm (+ SiLU)
# timm `BatchNormAct2d` / torchvision `Conv2dNormActivation`: F.batch_norm(training=True) then
# F.silu.  Pinned against those two torch functions and their autograd (tests/test_host_logic.py).
def _sigmoid(u):
    return 1.0 / (1.0 + np.exp(-u))


def bn_act_fwd_ref(x: np.ndarray, w: np.ndarray, b: np.ndarray, eps: float, silu: bool):
    """x [M,C] -> (y [M,C], mean [C], biased var [C]); statistics over the M rows."""
    x = np.asarray(x, np.float64)
    mean = x.mean(axis=0)
    var = ((x - mean) ** 2).mean(axis=0)
    u = (x - mean) / np.sqrt(var + eps) * np.asarray(w, np.float64) + np.asarray(b, np.float64)
    return (u * _sigmoid(u) if silu else u), mean, var


def bn_running_update_ref(running_mean, running_var, mean, var, M: int, momentum: float):
    """nn.BatchNorm2d's buffer update: unbiased variance, exponential average."""
    rm = (1 - momentum) * np.asarray(running_mean, np.float64) + momentum * mean
    rv = (1 - momentum) * np.asarray(running_var, np.float64) + momentum * var * M / max(M - 1, 1)
    return rm, rv


def bn_act_bwd_ref(g: np.ndarray, x: np.ndarray, w: np.ndarray, b: np.ndarray, eps: float, silu: bool):
    """-> (dx [M,C], dw [C], db [C]) of y = act(bn(x)) with batch statistics."""
    g = np.asarray(g, np.float64)
    x = np.asarray(x, np.float64)
    w = np.asarray(w, np.float64)
    M = x.shape[0]
    mean = x.mean(axis=0)
    rstd = 1.0 / np.sqrt(((x - mean) ** 2).mean(axis=0) + eps)
    xh = (x - mean) * rstd
    if silu:
        u = xh * w + np.asarray(b, np.float64)
        s = _sigmoid(u)
        du = g * s * (1.0 + u * (1.0 - s))
    else:
        du = g
    db = du.sum(axis=0)
    dw = (du * xh).sum(axis=0)
    dx = w * rstd * (du - db / M - xh * dw / M)
    return dx, dw, db


# ----------------------------------------------------------------------- squeeze-excite
def se_fwd_ref(x: np.ndarray, w1, b1, w2, b2):
    """x [B,P,C]; w1 [Cs,C], w2 [C,Cs] -> (y, gate [B,C]): x * sigmoid(fc2(silu(fc1(mean_p x))))."""
    x = np.asarray(x, np.float64)
    pooled = x.mean(axis=1)
    h = pooled @ np.asarray(w1, np.float64).T + np.asarray(b1, np.float64)
    a = h * _sigmoid(h)
    gate = _sigmoid(a @ np.asarray(w2, np.float64).T + np.asarray(b2, np.float64))
    return x * gate[:, None, :], gate


def se_bwd_ref(g: np.ndarray, x: np.ndarray, w1, b1, w2, b2):
    """-> (dx [B,P,C], dw1, db1, dw2, db2)."""
    g = np.asarray(g, np.float64)
    x = np.asarray(x, np.float64)
    w1 = np.asarray(w1, np.float64)
    w2 = np.asarray(w2, np.float64)
    P = x.shape[1]
    pooled = x.mean(axis=1)
    h = pooled @ w1.T + np.asarray(b1, np.float64)
    sh = _sigmoid(h)
    a = h * sh
    gate = _sigmoid(a @ w2.T + np.asarray(b2, np.float64))
    dgate = (g * x).sum(axis=1)
    dz2 = dgate * gate * (1.0 - gate)
    dw2, db2 = dz2.T @ a, dz2.sum(axis=0)
    dh = (dz2 @ w2) * sh * (1.0 + h * (1.0 - sh))
    dw1, db1 = dh.T @ pooled, dh.sum(axis=0)
    dpool = dh @ w1
    dx = g * gate[:, None, :] + dpool[:, None, :] / P
    return dx, dw1, db1, dw2, db2
