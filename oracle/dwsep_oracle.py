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
