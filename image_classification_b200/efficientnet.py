"""Depthwise-separable convolutions of EfficientNet on libcnx_b200 (BASELINE
config 5).

The older version of the reference trains timm `efficientnet_b2`
(/root/reference/previous/V3-efb2/config.py:10-15, model built at train_advanced_v3.0.py:330, AMP
loop :72-153).  Its MBConv blocks are  1x1 expand conv -> BN -> SiLU -> depthwise 3x3 / 5x5
(stride 1 / 2) -> BN -> SiLU -> squeeze-excite -> 1x1 project conv -> BN (+ residual).  This
module moves the two convolution kinds onto the library:

* `DepthwiseConv2dB200`  — `nn.Conv2d(C, C, k, stride, k//2, groups=C, bias=False)` ->
  `cnx_dwconv_kxk_fwd / _bwd` (NHWC bf16, fp32 accumulation);
* `PointwiseConv2dB200`  — `nn.Conv2d(Cin, Cout, 1)` on a [B,Cin,H,W] map -> the tcgen05 GEMM
  (`cnx_gemm_bf16`; data / weight gradients are the same kernel, the bias gradient a column sum).

Both are `nn.Conv2d` subclasses holding the SAME Parameter objects after
`swap_depthwise_separable(model)`, so `state_dict()`, the optimizer and SWA / EMA copies are
untouched; inputs that do not fit the fast path (CPU tensors, no bf16 autocast, odd shapes, the
[B,C,1,1] squeeze-excite convolutions) take `nn.Conv2d.forward`.  BatchNorm, SiLU and squeeze-excite
stay PyTorch ops in this first slice of f-4.  timm is not installable here; the
swap is written against `nn.Conv2d` attributes only, so it applies to timm's and torchvision's
EfficientNets alike (the parity tests use torchvision's `efficientnet_b2`, same architecture).
"""
from __future__ import annotations

import torch
from torch import nn

from . import ops


def _autocast_bf16() -> bool:
    return torch.is_autocast_enabled() and torch.get_autocast_dtype("cuda") == torch.bfloat16


def _nhwc_bf16(x: torch.Tensor) -> torch.Tensor:
    """[B,C,H,W] any float dtype / layout -> [B,H,W,C] bf16 contiguous (a view when the stream is
    already bf16 channels_last, which it is between the layers of a swapped model)."""
    if x.dtype != torch.bfloat16:
        x = x.to(torch.bfloat16)
    xn = x.permute(0, 2, 3, 1)
    return xn if xn.is_contiguous() else xn.contiguous()


class _DepthwiseFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, k, stride):
        C = weight.shape[0]
        xn = _nhwc_bf16(x)
        wt = weight.detach().reshape(C, k * k).t().contiguous().float()       # [k*k, C] tap-major
        y = ops.dwconv_kxk_fwd(xn, wt, k, stride)
        ctx.save_for_backward(xn, wt)
        ctx.k, ctx.stride, ctx.x_dtype = k, stride, x.dtype
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, g):
        xn, wt = ctx.saved_tensors
        k, C = ctx.k, wt.shape[1]
        dx, dwt = ops.dwconv_kxk_bwd(_nhwc_bf16(g), xn, wt, k, ctx.stride,
                                     need_dx=ctx.needs_input_grad[0], need_dw=ctx.needs_input_grad[1])
        if dx is not None:
            dx = dx.permute(0, 3, 1, 2).to(ctx.x_dtype)
        dw = dwt.t().reshape(C, 1, k, k).contiguous() if dwt is not None else None
        return dx, dw, None, None


class DepthwiseConv2dB200(nn.Conv2d):
    """Depthwise `nn.Conv2d` (k in {3,5}, stride in {1,2}, padding k//2, no bias) whose CUDA path
    under bf16 autocast (or on bf16 activations) is `cnx_dwconv_kxk_fwd/bwd`."""

    def _fast(self, x: torch.Tensor) -> bool:
        k, s = self.kernel_size[0], self.stride[0]
        return (x.is_cuda and x.dim() == 4 and self.bias is None
                and self.groups == self.in_channels == self.out_channels
                and self.kernel_size == (k, k) and k in (3, 5) and self.stride == (s, s) and s in (1, 2)
                and self.padding == (k // 2, k // 2) and self.dilation == (1, 1)
                and self.padding_mode == "zeros" and self.in_channels % 8 == 0
                and self.weight.dtype == torch.float32
                and (x.dtype == torch.bfloat16 or (x.is_floating_point() and _autocast_bf16())))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self._fast(x):
            return _DepthwiseFn.apply(x, self.weight, self.kernel_size[0], self.stride[0])
        return super().forward(x)


class _PointwiseFn(torch.autograd.Function):
    """z[M, Cout] = P[M, Cin] W^T (+ b) on the tcgen05 GEMM; dW = g^T P (split-K fp32),
    db = colsum(g), dP = g W."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        B, Cin, H, W = x.shape
        Cout = weight.shape[0]
        P = _nhwc_bf16(x).view(B * H * W, Cin)
        wq = weight.detach().reshape(Cout, Cin).to(torch.bfloat16)
        if bias is not None:
            z = ops.gemm_bf16(P, wq, epilogue=ops.EPI_BIAS, bias=bias.detach().float())
        else:
            z = ops.gemm_bf16(P, wq)
        ctx.save_for_backward(P, wq)
        ctx.shape, ctx.x_dtype, ctx.has_bias = (B, H, W), x.dtype, bias is not None
        return z.view(B, H, W, Cout).permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, g):
        P, wq = ctx.saved_tensors
        B, H, W = ctx.shape
        Cout, Cin = wq.shape
        g2 = _nhwc_bf16(g).view(B * H * W, Cout)
        dw = None
        if ctx.needs_input_grad[1]:
            dw = ops.gemm_bf16(g2, P, a_mn_major=True, b_mn_major=True,
                               epilogue=ops.EPI_STORE_F32).view(Cout, Cin, 1, 1)
        db = ops.colsum_bf16(g2) if ctx.has_bias and ctx.needs_input_grad[2] else None
        dx = None
        if ctx.needs_input_grad[0]:
            dP = ops.gemm_bf16(g2, wq, b_mn_major=True)
            dx = dP.view(B, H, W, Cin).permute(0, 3, 1, 2).to(ctx.x_dtype)
        return dx, dw, db


class PointwiseConv2dB200(nn.Conv2d):
    """1x1 `nn.Conv2d` (stride 1, no padding, groups 1) whose CUDA path under bf16 autocast is the
    library GEMM.  Maps with fewer than `min_rows` pixels in the batch (the squeeze-excite
    convolutions on [B,C,1,1]) stay on `nn.Conv2d.forward`: a 128-row tile would be mostly empty."""

    min_rows = 1024

    def _fast(self, x: torch.Tensor) -> bool:
        return (x.is_cuda and x.dim() == 4 and self.kernel_size == (1, 1) and self.stride == (1, 1)
                and self.padding == (0, 0) and self.dilation == (1, 1) and self.groups == 1
                and self.in_channels % 8 == 0 and self.out_channels % 8 == 0
                and self.weight.dtype == torch.float32
                and x.shape[0] * x.shape[2] * x.shape[3] >= self.min_rows
                and (x.dtype == torch.bfloat16 or x.is_floating_point()) and _autocast_bf16())

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self._fast(x):
            return _PointwiseFn.apply(x, self.weight, self.bias)
        return super().forward(x)


def _rebind(conv: nn.Conv2d, cls) -> nn.Conv2d:
    new = cls(conv.in_channels, conv.out_channels, conv.kernel_size, conv.stride, conv.padding,
              conv.dilation, conv.groups, conv.bias is not None, conv.padding_mode)
    new.weight = conv.weight                  # the same Parameter objects
    if conv.bias is not None:
        new.bias = conv.bias
    return new.train(conv.training)


def swap_depthwise_separable(model: nn.Module) -> tuple[int, int]:
    """Replace, anywhere in `model`, depthwise 3x3 / 5x5 convolutions by `DepthwiseConv2dB200` and
    1x1 convolutions by `PointwiseConv2dB200`, around the same Parameter objects.  Returns
    (depthwise replaced, pointwise replaced)."""
    n_dw = n_pw = 0
    for parent in list(model.modules()):
        for name, child in list(parent.named_children()):
            if type(child) is not nn.Conv2d:
                continue
            k = child.kernel_size
            if (child.groups == child.in_channels == child.out_channels and child.groups > 1
                    and k in ((3, 3), (5, 5)) and child.bias is None
                    and child.padding == (k[0] // 2, k[0] // 2) and child.in_channels % 8 == 0):
                setattr(parent, name, _rebind(child, DepthwiseConv2dB200))
                n_dw += 1
            elif (k == (1, 1) and child.groups == 1 and child.stride == (1, 1)
                  and child.padding == (0, 0) and child.in_channels % 8 == 0
                  and child.out_channels % 8 == 0):
                setattr(parent, name, _rebind(child, PointwiseConv2dB200))
                n_pw += 1
    return n_dw, n_pw
