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
[B,C,1,1] squeeze-excite convolutions) take `nn.Conv2d.forward`.  The memory-bound glue between
the convolutions is on the library too (csrc/bnse.cu):

* `BatchNormActB200`     — `nn.BatchNorm2d` (+ the SiLU that follows it) -> `cnx_bn_stats` +
  `cnx_bn_act_fwd`, backward `cnx_bn_act_bwd`; running statistics updated like nn.BatchNorm2d;
* `SqueezeExciteB200`    — x * sigmoid(fc2(silu(fc1(mean x)))): the pooled sums come out of the
  preceding `BatchNormActB200` pass, the two [B,C] gate layers are `cnx_se_gate_fwd/bwd` (fp32),
  the scaling and its backward are `cnx_se_scale_*`;
* `swap_efficientnet(model)` installs all four kinds.

timm is not installable here; the swap recognises modules by structure (torchvision's
conv -> BatchNorm2d -> SiLU sequences and `SqueezeExcitation`, timm's `BatchNormAct2d` and
`SqueezeExcite` layouts), so it applies to both libraries' EfficientNets: the parity tests use
torchvision's `efficientnet_b2` (same architecture) and timm-layout stand-ins (tests/_foreign.py).
A CUDA activation that misses the library path because autocast is off or not bf16 (the
reference's own default is fp16, train_advanced_v3.0.py:103) is reported once with a warning —
the modules then behave like the torch modules they subclass, they never route through a CPU or
oracle path, and on the library path a missing `libcnx_b200.so` raises.
"""
from __future__ import annotations

import torch
from torch import nn

from . import _lib, ops


_warned: set[str] = set()


def _note_off_path(who: str, x: torch.Tensor) -> None:
    """One warning per module kind when a CUDA activation runs on torch's kernels instead."""
    if x.is_cuda and who not in _warned:
        _warned.add(who)
        import warnings
        warnings.warn(f"{who}: CUDA input of dtype {x.dtype} is not on the libcnx_b200 path "
                      "(needs bf16 activations or torch.autocast('cuda', dtype=torch.bfloat16)); "
                      "running the torch module it subclasses", stacklevel=3)


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
        if x.dtype != torch.bfloat16 and not _autocast_bf16():
            _note_off_path("DepthwiseConv2dB200", x)
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
        if not _autocast_bf16():
            _note_off_path("PointwiseConv2dB200", x)
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


# ----------------------------------------------------------------------- BatchNorm (+ SiLU)
class _BnActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, training, momentum, eps, act,
                want_pool):
        xn = _nhwc_bf16(x)
        w, b = weight.detach().float().contiguous(), bias.detach().float().contiguous()
        # the result is allocated as the [B,C,H,W] channels_last tensor the caller gets (not a view
        # made inside this Function: torchvision's MBConv adds the shortcut to it in place)
        B, H, W, C = xn.shape
        y = torch.empty((B, C, H, W), device=xn.device, dtype=torch.bfloat16,
                        memory_format=torch.channels_last)
        _, mean, rstd, pool = ops.bn_act_fwd(xn, w, b, running_mean, running_var, training=training,
                                             momentum=momentum, eps=eps, act=act, want_pool=want_pool,
                                             out=y.permute(0, 2, 3, 1))
        ctx.save_for_backward(xn, w, b, mean, rstd)
        ctx.act, ctx.training, ctx.x_dtype = act, training, x.dtype
        if pool is None:
            pool = xn.new_empty(0)
        ctx.mark_non_differentiable(pool)
        return y, pool

    @staticmethod
    def backward(ctx, g, _gpool):
        if not ctx.training:
            raise _lib.CnxError("BatchNormActB200: backward through eval-mode statistics is not "
                                "implemented (the reference never trains in eval mode)")
        xn, w, b, mean, rstd = ctx.saved_tensors
        dx, dw, db = ops.bn_act_bwd(_nhwc_bf16(g), xn, w, b, mean, rstd, ctx.act)
        dx = dx.permute(0, 3, 1, 2)
        if ctx.x_dtype != torch.bfloat16:
            dx = dx.to(ctx.x_dtype)
        return dx, dw, db, None, None, None, None, None, None, None


class BatchNormActB200(nn.BatchNorm2d):
    """`nn.BatchNorm2d` followed by `act` ("silu" or None) as ONE pass over the NHWC bf16 map
    (plus the statistics pass in training).  `emit_pool`: also leave the per-image channel sums of
    the output on the returned tensor (`_cnx_pool_sum`) for the squeeze-excite block that follows."""

    act: str | None = None
    emit_pool: bool = False

    def _fast(self, x: torch.Tensor) -> bool:
        return (x.is_cuda and x.dim() == 4 and self.affine and self.num_features % 8 == 0
                and self.momentum is not None and self.weight.dtype == torch.float32
                and (self.training or self.running_mean is not None)
                and (x.dtype == torch.bfloat16 or (x.is_floating_point() and _autocast_bf16())))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not self._fast(x):
            if x.dtype != torch.bfloat16 and not _autocast_bf16():
                _note_off_path("BatchNormActB200", x)
            y = super().forward(x)
            return torch.nn.functional.silu(y) if self.act == "silu" else y
        training = self.training or self.running_mean is None
        if training and self.num_batches_tracked is not None:
            self.num_batches_tracked.add_(1)
        y, pool = _BnActFn.apply(x, self.weight, self.bias, self.running_mean, self.running_var,
                                 training, float(self.momentum), float(self.eps),
                                 ops.ACT_SILU if self.act == "silu" else ops.ACT_NONE,
                                 self.emit_pool)
        if self.emit_pool:
            y._cnx_pool_sum = pool
        return y


# ----------------------------------------------------------------------- squeeze-excite
class _SqueezeExciteFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, pool_sum, w1, b1, w2, b2):
        xn = _nhwc_bf16(x)
        B, H, W, C = xn.shape
        if pool_sum is not None and pool_sum.numel() == B * C:
            pool, scale = pool_sum, 1.0 / (H * W)          # sums handed over by the BatchNorm pass
        else:
            pool, scale = ops.avgpool_nhwc_fwd(xn).float(), 1.0
        Cs = w1.shape[0]
        w1m = w1.detach().float().reshape(Cs, C).contiguous()
        w2t = w2.detach().float().reshape(C, Cs).t().contiguous()         # [Cs, C]
        hidden, gate = ops.se_gate_fwd(pool, scale, w1m, b1.detach().float().contiguous(), w2t,
                                       b2.detach().float().contiguous())
        y = ops.se_scale_fwd(xn, gate)
        ctx.save_for_backward(xn, pool, gate, hidden, w1m, w2t)
        ctx.scale, ctx.x_dtype, ctx.w_shapes = scale, x.dtype, (w1.shape, w2.shape)
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, g):
        xn, pool, gate, hidden, w1m, w2t = ctx.saved_tensors
        gn = _nhwc_bf16(g)
        dgate = ops.se_scale_bwd_reduce(gn, xn)
        dpool, dw1, db1, dw2t, db2 = ops.se_gate_bwd(dgate, gate, hidden, pool, ctx.scale, w1m, w2t)
        dx = ops.se_scale_bwd_apply(gn, gate, dpool).permute(0, 3, 1, 2)
        if ctx.x_dtype != torch.bfloat16:
            dx = dx.to(ctx.x_dtype)
        return (dx, None, dw1.reshape(ctx.w_shapes[0]), db1,
                dw2t.t().contiguous().reshape(ctx.w_shapes[1]), db2)


class SqueezeExciteB200(nn.Module):
    """Drop-in for torchvision `SqueezeExcitation` / timm `SqueezeExcite` with SiLU + sigmoid:
    holds the SAME two 1x1 convolution modules under their original attribute names."""

    def __init__(self, reduce_conv: nn.Conv2d, expand_conv: nn.Conv2d,
                 names: tuple[str, str] = ("fc1", "fc2")):
        super().__init__()
        self._names = names
        setattr(self, names[0], reduce_conv)
        setattr(self, names[1], expand_conv)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        r, e = getattr(self, self._names[0]), getattr(self, self._names[1])
        fast = (x.is_cuda and x.dim() == 4 and x.shape[1] % 8 == 0
                and (x.dtype == torch.bfloat16 or (x.is_floating_point() and _autocast_bf16())))
        if not fast:
            s = torch.sigmoid(e(torch.nn.functional.silu(r(x.mean((2, 3), keepdim=True)))))
            return s * x
        return _SqueezeExciteFn.apply(x, getattr(x, "_cnx_pool_sum", None), r.weight, r.bias,
                                      e.weight, e.bias)


def _rebind_bn(bn: nn.BatchNorm2d, act: str | None) -> BatchNormActB200:
    new = BatchNormActB200(bn.num_features, bn.eps, bn.momentum, bn.affine, bn.track_running_stats)
    if bn.affine:
        new.weight, new.bias = bn.weight, bn.bias                 # the same Parameter objects
    if bn.track_running_stats:
        new.running_mean, new.running_var = bn.running_mean, bn.running_var
        new.num_batches_tracked = bn.num_batches_tracked
    new.act = act
    return new.train(bn.training)


def _se_names(m: nn.Module):
    """(reduce, expand) attribute names if `m` is a SiLU / sigmoid squeeze-excite block in
    torchvision's (`fc1`, `fc2`, `activation`, `scale_activation`) or timm's (`conv_reduce`,
    `conv_expand`, `act1`, `gate`) layout, else None."""
    for red, exp, act, gate in (("fc1", "fc2", "activation", "scale_activation"),
                                ("conv_reduce", "conv_expand", "act1", "gate")):
        if (all(hasattr(m, a) for a in (red, exp, act, gate)) and not isinstance(m, SqueezeExciteB200)
                and isinstance(getattr(m, act), nn.SiLU) and isinstance(getattr(m, gate), nn.Sigmoid)
                and isinstance(getattr(m, red), nn.Conv2d) and isinstance(getattr(m, exp), nn.Conv2d)
                and getattr(m, red).bias is not None and getattr(m, exp).bias is not None):
            return red, exp
    return None


def _fused_bn_act(m: nn.Module):
    """timm's `BatchNormAct2d` layout: an nn.BatchNorm2d subclass carrying its activation in
    `.act` (SiLU or Identity) and a no-op `.drop`.  Returns (True, act) when it can be rebound."""
    if (isinstance(m, nn.BatchNorm2d) and type(m) is not nn.BatchNorm2d
            and not isinstance(m, BatchNormActB200) and hasattr(m, "act")
            and isinstance(getattr(m, "drop", nn.Identity()), nn.Identity)):
        if isinstance(m.act, nn.SiLU):
            return True, "silu"
        if isinstance(m.act, nn.Identity):
            return True, None
    return False, None


def swap_efficientnet(model: nn.Module) -> dict:
    """Install the library modules in an EfficientNet: depthwise / 1x1 convolutions
    (`swap_depthwise_separable`); every `nn.BatchNorm2d` inside an `nn.Sequential` (torchvision's
    conv -> BatchNorm2d -> SiLU layout: the `nn.SiLU` after it is absorbed and replaced by
    `nn.Identity`) and every BatchNorm-with-activation module of timm's `BatchNormAct2d` layout;
    every SiLU / sigmoid squeeze-excite block in either library's naming.  The BatchNorm right
    before a squeeze-excite block also emits its pooled sums.  Parameters and buffers are the same
    objects, state_dict keys do not change.  Returns the counts."""
    n_dw, n_pw = swap_depthwise_separable(model)
    n_bn = n_se = 0
    for parent in list(model.modules()):
        kids = list(parent.named_children())
        for i, (name, child) in enumerate(kids):
            if child.__class__ is nn.BatchNorm2d and isinstance(parent, nn.Sequential):
                if not (child.affine and child.num_features % 8 == 0):
                    continue
                nxt = kids[i + 1] if i + 1 < len(kids) else None
                act = "silu" if nxt is not None and type(nxt[1]) is nn.SiLU else None
                setattr(parent, name, _rebind_bn(child, act))
                if act:
                    setattr(parent, nxt[0], nn.Identity())
                n_bn += 1
            else:
                ok, act = _fused_bn_act(child)
                if ok and child.affine and child.num_features % 8 == 0:
                    setattr(parent, name, _rebind_bn(child, act))
                    n_bn += 1
    for parent in list(model.modules()):
        kids = list(parent.named_children())
        for i, (name, child) in enumerate(kids):
            names = _se_names(child)
            if names is None:
                continue
            se = SqueezeExciteB200(getattr(child, names[0]), getattr(child, names[1]), names)
            setattr(parent, name, se.train(child.training))
            n_se += 1
            if i > 0:                              # the module whose output this block pools
                prev = getattr(parent, kids[i - 1][0])      # as rebound by the loop above
                if isinstance(prev, nn.Sequential):
                    tail = [m for m in prev.children() if not isinstance(m, nn.Identity)]
                    prev = tail[-1] if tail else None
                if isinstance(prev, BatchNormActB200):
                    prev.emit_pool = True
    return {"depthwise": n_dw, "pointwise": n_pw, "batchnorm": n_bn, "squeeze_excite": n_se}
