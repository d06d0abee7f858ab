"""A timm-compatible ConvNeXt skeleton to host the block.

`timm` is not installable here, so the object `timm.create_model('convnext_base',
num_classes=44)` returns (/root/reference/train_advanced_v4.py:488-492) is re-stated with the
same attribute layout the reference touches: `stem`, `stages[i].downsample`, `stages[i].blocks[j]`,
`head.{global_pool,norm,flatten,drop,fc}` (:106-111, :131-143).  The 36 blocks, the stem, the
stage downsamples and the head's pool + norm run on libcnx_b200 kernels; only
the 44-way `head.fc` / auxiliary Linear layers on [B, C] stay `nn.Linear`.  A CUDA activation that
misses a library path warns once (and raises under CNX_STRICT=1) — see `_off_path`.  The residual stream is
kept channels_last so the NHWC kernels are zero-copy between blocks.  `DeepSupervisionModelB200`
mirrors the wrapper the reference puts around the backbone (:96-152).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn

import os
import warnings

from . import _lib, ops
from .block import ConvNeXtBlockB200

_warned: set[str] = set()


def _off_path(who: str, why: str, x: torch.Tensor) -> None:
    """A CUDA activation that misses the libcnx_b200 path of a trunk module is never silent
    ("unsupported shape/dtype => error, not a silent fallback"): under
    CNX_STRICT=1 (bench.py and the GPU tests run that way) it raises; otherwise it warns once per
    module kind, naming the failed condition, and the torch module it subclasses runs.  CPU
    tensors (the oracle side of the tests) take the torch ops without comment."""
    if not x.is_cuda:
        return
    msg = (f"{who}: CUDA input (dtype {x.dtype}, shape {tuple(x.shape)}) is not on the libcnx_b200 "
           f"path: {why}")
    if os.environ.get("CNX_STRICT") == "1":
        raise _lib.CnxError(msg + " (CNX_STRICT=1)")
    if who not in _warned:
        _warned.add(who)
        warnings.warn(msg + "; running the torch module it subclasses", stacklevel=3)


def _first_failed(conds) -> str | None:
    """conds: iterable of (ok, description); -> description of the first failed one or None."""
    for ok, why in conds:
        if not ok:
            return why
    return None

CONVNEXT_CFGS = {
    "convnext_tiny": dict(depths=(3, 3, 9, 3), dims=(96, 192, 384, 768)),
    "convnext_small": dict(depths=(3, 3, 27, 3), dims=(96, 192, 384, 768)),
    "convnext_base": dict(depths=(3, 3, 27, 3), dims=(128, 256, 512, 1024)),
    "convnext_large": dict(depths=(3, 3, 27, 3), dims=(192, 384, 768, 1536)),
}


class _LayerNorm2dFn(torch.autograd.Function):
    """x: [B,C,H,W] bf16 channels_last (physically NHWC rows) -> same layout, libcnx_b200 kernels."""

    @staticmethod
    def forward(ctx, x, weight, bias, eps):
        B, C, H, W = x.shape
        x2d = x.permute(0, 2, 3, 1).reshape(B * H * W, C)
        y, mean, rstd = ops.layernorm_fwd(x2d, weight.detach(), bias.detach(), eps)
        ctx.save_for_backward(x2d, weight, mean, rstd)
        return y.view(B, H, W, C).permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, g):
        x2d, weight, mean, rstd = ctx.saved_tensors
        B, C, H, W = g.shape
        g = g.permute(0, 2, 3, 1)
        if g.dtype != torch.bfloat16 or not g.is_contiguous():
            g = g.to(torch.bfloat16).contiguous()
        dx, dw, db = ops.layernorm_bwd(g.view(B * H * W, C), x2d, weight.detach(), mean, rstd)
        return dx.view(B, H, W, C).permute(0, 3, 1, 2), dw, db, None


class LayerNorm2d(nn.LayerNorm):
    """LayerNorm over the channel dim of an NCHW tensor (timm's LayerNorm2d).  bf16 channels_last
    CUDA activations (the stem / downsample stream of the B200 model) go through
    cnx_layernorm_fwd/bwd (the pooled [B,C,1,1] map of the head included); anything else (the
    oracle on CPU) takes the PyTorch op."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        why = _first_failed((
            (x.is_cuda, "not a CUDA tensor"),
            (x.dim() == 4, "input is not [B,C,H,W]"),
            (x.dtype == torch.bfloat16, "activations are not bf16 (run the model under "
                                        "torch.autocast('cuda', dtype=torch.bfloat16))"),
            (x.dim() == 4 and x.shape[1] % 8 == 0 and x.shape[1] <= 1024,
             "channel count must be a multiple of 8 and <= 1024"),
            (x.dim() == 4 and x.permute(0, 2, 3, 1).is_contiguous(), "input is not channels_last"),
            (self.weight.dtype == torch.float32, "parameters are not float32"),
        ))
        if why is None:
            return _LayerNorm2dFn.apply(x, self.weight, self.bias, self.eps)
        _off_path("LayerNorm2d", why, x)
        y = x.permute(0, 2, 3, 1)
        y = F.layer_norm(y, self.normalized_shape, self.weight, self.bias, self.eps)
        # autocast runs layer_norm in fp32; hand the stream back in the dtype it came in so the
        # blocks keep their zero-copy bf16 channels_last path
        return y.to(x.dtype).permute(0, 3, 1, 2)


class PatchConv2d(nn.Conv2d):
    """The stem (4x4 stride 4) and downsample (2x2 stride 2) convolutions of timm's ConvNeXt: a
    plain `nn.Conv2d` (same parameters, same state_dict keys) that marks the convolution the
    containers below (`StemB200`, `DownsampleB200`) run as a patchify + tcgen05 GEMM.  Called on
    its own it IS `nn.Conv2d` — there is no second library-backed implementation of this op."""


def _w2d(weight: torch.Tensor) -> torch.Tensor:
    """conv weight [Cout, Cin, k, k] -> [Cout, k*k*Cin] bf16 with inner order (kh, kw, c): a view
    + cast when the parameter is channels_last (what `create_convnext(...).to(channels_last)` and
    bench.py use), one small permute copy otherwise."""
    return weight.detach().permute(0, 2, 3, 1).reshape(weight.shape[0], -1).to(torch.bfloat16)


def _dw4d(dw2d: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
    """[Cout, k*k*Cin] fp32 gradient -> a tensor shaped and laid out like `weight`."""
    Cout, Cin, k, _ = weight.shape
    g = dw2d.view(Cout, k, k, Cin).permute(0, 3, 1, 2)
    if g.stride() != weight.stride():
        out = torch.empty_like(weight, dtype=torch.float32)
        out.copy_(g)
        return out
    return g


def _autocast_bf16() -> bool:
    return torch.is_autocast_enabled() and torch.get_autocast_dtype("cuda") == torch.bfloat16


class _PatchGemmFn(torch.autograd.Function):
    """k x k stride-k convolution + bias on a patch-major operand P [M, k*k*Cin] bf16 as GEMMs of
    libcnx_b200: z = P W^T + b (EPI_BIAS), dW = g^T P (split-K fp32), db = colsum(g),
    dP = g W (only when P needs a gradient)."""

    @staticmethod
    def forward(ctx, P, weight, bias):
        wq = _w2d(weight)
        ctx.save_for_backward(P, wq)
        ctx.w_like = weight.detach()
        return ops.gemm_bf16(P, wq, epilogue=ops.EPI_BIAS, bias=bias.detach().float())

    @staticmethod
    def backward(ctx, g):
        P, wq = ctx.saved_tensors
        g = g.contiguous()
        dw = ops.gemm_bf16(g, P, a_mn_major=True, b_mn_major=True, epilogue=ops.EPI_STORE_F32)
        db = ops.colsum_bf16(g)
        dP = ops.gemm_bf16(g, wq, b_mn_major=True) if ctx.needs_input_grad[0] else None
        return dP, _dw4d(dw, ctx.w_like), db


class _LayerNormPatchFn(torch.autograd.Function):
    """LayerNorm2d of x [B,C,H,W] bf16 channels_last whose output is the patch-major operand of the
    following k x k stride-k convolution ([B*H/k*W/k, k*k*C]); backward takes the patch-major
    gradient the data-gradient GEMM produced."""

    @staticmethod
    def forward(ctx, x, weight, bias, eps, patch):
        xn = x.permute(0, 2, 3, 1)
        y, mean, rstd = ops.layernorm_patch_fwd(xn, weight.detach(), bias.detach(), patch, eps)
        ctx.save_for_backward(xn, weight, mean, rstd)
        ctx.patch = patch
        return y

    @staticmethod
    def backward(ctx, g):
        xn, weight, mean, rstd = ctx.saved_tensors
        dx, dw, db = ops.layernorm_patch_bwd(g.contiguous(), xn, weight.detach(), mean, rstd, ctx.patch)
        return dx.permute(0, 3, 1, 2), dw, db, None, None


class DownsampleB200(nn.Sequential):
    """timm's stage `downsample` = Sequential(LayerNorm2d, Conv2d(k=2, s=2)) with the same children
    and state_dict keys.  bf16 channels_last CUDA activations under bf16 autocast run as
    cnx_layernorm_patch_fwd -> tcgen05 GEMM (+bias); anything else takes the children one by one."""

    def forward(self, x):
        norm, conv = self[0], self[1]
        k = conv.kernel_size[0]
        why = _first_failed((
            (x.is_cuda, "not a CUDA tensor"),
            (x.dim() == 4, "input is not [B,C,H,W]"),
            (x.dtype == torch.bfloat16, "activations are not bf16"),
            (_autocast_bf16(), "torch.autocast('cuda', dtype=torch.bfloat16) is not active"),
            (x.dim() == 4 and x.shape[1] % 8 == 0 and x.shape[1] <= 1024,
             "channel count must be a multiple of 8 and <= 1024"),
            (conv.kernel_size == (k, k) and conv.stride == (k, k) and conv.padding == (0, 0)
             and conv.groups == 1 and conv.bias is not None,
             "the convolution is not a k x k stride-k patchify convolution with bias"),
            (x.dim() == 4 and x.shape[2] % k == 0 and x.shape[3] % k == 0,
             "spatial size is not a multiple of the patch"),
            (conv.out_channels % 8 == 0, "output channels must be a multiple of 8"),
            (x.dim() == 4 and x.permute(0, 2, 3, 1).is_contiguous(), "input is not channels_last"),
            (norm.weight.dtype == torch.float32, "parameters are not float32"),
        ))
        if why is None:
            B, C, H, W = x.shape
            yp = _LayerNormPatchFn.apply(x, norm.weight, norm.bias, norm.eps, k)
            z = _PatchGemmFn.apply(yp, conv.weight, conv.bias)
            return z.view(B, H // k, W // k, conv.out_channels).permute(0, 3, 1, 2)
        _off_path("DownsampleB200", why, x)
        return conv(F.layer_norm(x.permute(0, 2, 3, 1), norm.normalized_shape, norm.weight,
                                 norm.bias, norm.eps).permute(0, 3, 1, 2))


class StemB200(nn.Sequential):
    """timm's `stem` = Sequential(Conv2d(k=4, s=4), LayerNorm2d).  A contiguous fp32 NCHW CUDA batch
    under bf16 autocast (what the loop feeds, /root/reference/train_advanced_v4.py:211-218) runs as
    cnx_patchify_nchw_f32 -> tcgen05 GEMM (+bias) -> cnx_layernorm_fwd."""

    def forward(self, x):
        conv, norm = self[0], self[1]
        k = conv.kernel_size[0]
        why = _first_failed((
            (x.is_cuda, "not a CUDA tensor"),
            (x.dim() == 4, "input is not [B,C,H,W]"),
            (x.dtype == torch.float32 and x.is_contiguous(),
             "the image batch is not contiguous float32 NCHW"),
            (_autocast_bf16(), "torch.autocast('cuda', dtype=torch.bfloat16) is not active"),
            (not x.requires_grad, "the image batch requires a gradient"),
            (conv.kernel_size == (k, k) and conv.stride == (k, k) and conv.padding == (0, 0)
             and conv.groups == 1 and conv.bias is not None,
             "the convolution is not a k x k stride-k patchify convolution with bias"),
            (x.dim() == 4 and x.shape[2] % k == 0 and x.shape[3] % k == 0,
             "spatial size is not a multiple of the patch"),
            (x.dim() == 4 and (k * k * x.shape[1]) % 8 == 0 and conv.out_channels % 8 == 0,
             "patch size / output channels must be multiples of 8"),
        ))
        if why is None:
            B, _, H, W = x.shape
            P = ops.patchify_nchw_f32(x, k)
            z = _PatchGemmFn.apply(P, conv.weight, conv.bias)
            return norm(z.view(B, H // k, W // k, conv.out_channels).permute(0, 3, 1, 2))
        _off_path("StemB200", why, x)
        y = conv(x)
        return F.layer_norm(y.permute(0, 2, 3, 1), norm.normalized_shape, norm.weight, norm.bias,
                            norm.eps).permute(0, 3, 1, 2)


class ConvNeXtStage(nn.Module):
    def __init__(self, in_chs: int, out_chs: int, depth: int, stride: int, block_cls, ls_init_value):
        super().__init__()
        if in_chs != out_chs or stride > 1:
            self.downsample = DownsampleB200(
                LayerNorm2d(in_chs, eps=1e-6),
                PatchConv2d(in_chs, out_chs, kernel_size=stride, stride=stride),
            )
        else:
            self.downsample = nn.Identity()
        self.blocks = nn.Sequential(*[block_cls(out_chs, ls_init_value=ls_init_value)
                                      for _ in range(depth)])

    def forward(self, x):
        return self.blocks(self.downsample(x))


class _GlobalAvgPoolFn(torch.autograd.Function):
    """[B,C,H,W] bf16 channels_last -> [B,C,1,1] bf16: cnx_avgpool_nhwc_fwd / _bwd."""

    @staticmethod
    def forward(ctx, x):
        ctx.hw = x.shape[2:]
        y = ops.avgpool_nhwc_fwd(x.permute(0, 2, 3, 1))
        return y.view(y.shape[0], y.shape[1], 1, 1)

    @staticmethod
    def backward(ctx, g):
        B, C = g.shape[:2]
        g2 = g.reshape(B, C).to(torch.bfloat16).contiguous()
        return ops.avgpool_nhwc_bwd(g2, *ctx.hw).permute(0, 3, 1, 2)


class GlobalAvgPoolB200(nn.AdaptiveAvgPool2d):
    """`nn.AdaptiveAvgPool2d(1)` (head.global_pool, /root/reference/train_advanced_v4.py:140, and the
    auxiliary classifiers, :119-121) reading the bf16 channels_last stream in place; any other
    input takes the PyTorch op."""

    def __init__(self, output_size=1):
        super().__init__(output_size)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        why = _first_failed((
            (x.is_cuda, "not a CUDA tensor"),
            (self.output_size in (1, (1, 1)), "output size is not 1"),
            (x.dim() == 4, "input is not [B,C,H,W]"),
            (x.dtype == torch.bfloat16, "activations are not bf16"),
            (x.dim() == 4 and x.shape[1] % 8 == 0 and x.shape[0] <= 65535,
             "channel count must be a multiple of 8 and the batch <= 65535"),
            (x.dim() == 4 and x.permute(0, 2, 3, 1).is_contiguous(), "input is not channels_last"),
        ))
        if why is None:
            return _GlobalAvgPoolFn.apply(x)
        _off_path("GlobalAvgPoolB200", why, x)
        return super().forward(x)


class _Head(nn.Module):
    def __init__(self, dim: int, num_classes: int):
        super().__init__()
        self.global_pool = GlobalAvgPoolB200(1)
        self.norm = LayerNorm2d(dim, eps=1e-6)
        self.flatten = nn.Flatten(1)
        self.drop = nn.Dropout(0.0)
        self.fc = nn.Linear(dim, num_classes)

    def forward(self, x):
        return self.fc(self.drop(self.flatten(self.norm(self.global_pool(x)))))


class ConvNeXtTimmLike(nn.Module):
    """stem -> 4 stages -> head, attribute-compatible with timm's ConvNeXt."""

    def __init__(self, depths=(3, 3, 27, 3), dims=(128, 256, 512, 1024), num_classes: int = 44,
                 in_chans: int = 3, ls_init_value: float = 1e-6, block_cls=ConvNeXtBlockB200):
        super().__init__()
        self.num_classes = num_classes
        self.stem = StemB200(PatchConv2d(in_chans, dims[0], kernel_size=4, stride=4),
                                  LayerNorm2d(dims[0], eps=1e-6))
        stages, prev = [], dims[0]
        for i, (d, c) in enumerate(zip(depths, dims)):
            stages.append(ConvNeXtStage(prev, c, d, 1 if i == 0 else 2, block_cls, ls_init_value))
            prev = c
        self.stages = nn.Sequential(*stages)
        self.num_features = prev
        self.head = _Head(prev, num_classes)
        self.apply(self._init_weights)

    @staticmethod
    def _init_weights(m):
        # trunc-normal(0.02) weights, zero bias: timm's and torchvision's ConvNeXt init
        if isinstance(m, (nn.Conv2d, nn.Linear)):
            nn.init.trunc_normal_(m.weight, std=0.02)
            if m.bias is not None:
                nn.init.zeros_(m.bias)

    def forward_features(self, x):
        return self.stages(self.stem(x))

    def forward(self, x):
        return self.head(self.forward_features(x))


class DeepSupervisionModelB200(nn.Module):
    """The wrapper `create_model` puts around the backbone when `cfg.use_deep_supervision` is set
    (config.py:22; /root/reference/train_advanced_v4.py:96-152, :494-495): same attributes
    (`base_model`, `aux_classifiers`), same forward (stem, stages, then head.global_pool / norm /
    flatten / fc child by child — `head.drop` is skipped there too) and the same return value
    `[main_out] + aux_outs`.  The training loop keeps `outputs[0]` only (:219-220), so the three
    auxiliary heads never receive a gradient; `compute_aux=False` skips their
    forward and returns `[main_out]`."""

    def __init__(self, base_model: nn.Module, num_classes: int, compute_aux: bool = True):
        super().__init__()
        self.base_model = base_model
        if not hasattr(base_model, "stages"):
            raise NotImplementedError("Deep supervision only implemented for ConvNeXt")   # :115
        channels = []
        for stage in base_model.stages[1:]:
            last_block = stage.blocks[-1]
            if hasattr(last_block, "conv_dw"):
                channels.append(last_block.conv_dw.out_channels)
            elif hasattr(last_block, "dwconv"):
                channels.append(last_block.dwconv.out_channels)
            else:
                raise NotImplementedError("cannot read the block's channel count")         # :113
        self.aux_classifiers = nn.ModuleList([
            nn.Sequential(GlobalAvgPoolB200(1), nn.Flatten(), nn.Linear(dim, num_classes))
            for dim in channels])
        self.compute_aux = compute_aux

    def forward(self, x):
        features = []
        x = self.base_model.stem(x)
        for i, stage in enumerate(self.base_model.stages):
            x = stage(x)
            if i > 0:
                features.append(x)
        head = self.base_model.head
        main_out = head.fc(head.flatten(head.norm(head.global_pool(x))))
        if not self.compute_aux:
            return [main_out]
        return [main_out] + [clf(f) for f, clf in zip(features, self.aux_classifiers)]


def _as_layernorm2d(m: nn.Module) -> nn.Module:
    if isinstance(m, LayerNorm2d) or not isinstance(m, nn.LayerNorm) or not m.elementwise_affine:
        return m
    new = LayerNorm2d(m.normalized_shape[0], eps=m.eps)
    new.weight, new.bias = m.weight, m.bias          # the same Parameter objects
    return new.train(m.training)


def _as_patch_conv(c: nn.Module) -> nn.Module:
    if isinstance(c, PatchConv2d) or not isinstance(c, nn.Conv2d) or c.bias is None:
        return c
    if c.kernel_size != c.stride or c.padding != (0, 0) or c.groups != 1 or c.dilation != (1, 1):
        return c
    new = PatchConv2d(c.in_channels, c.out_channels, kernel_size=c.kernel_size, stride=c.stride)
    new.weight, new.bias = c.weight, c.bias
    return new.train(c.training)


def swap_trunk(model: nn.Module) -> int:
    """The f-1 counterpart of `block.swap_blocks` for a timm-style ConvNeXt built elsewhere
    (`timm.create_model('convnext_base', ...)`, /root/reference/train_advanced_v4.py:488-492):
    `stem` -> StemB200, `stages[i].downsample` -> DownsampleB200, `head.global_pool` ->
    GlobalAvgPoolB200, `head.norm` -> this module's LayerNorm2d.  The new modules hold the SAME
    Parameter objects under the same names, so `state_dict()`, the optimizer and the EMA shadow are
    untouched.  Anything that does not look like timm's layout is left alone.  Returns the number
    of modules replaced."""
    n = 0
    stem = getattr(model, "stem", None)
    if (isinstance(stem, nn.Sequential) and not isinstance(stem, StemB200) and len(stem) == 2
            and isinstance(stem[0], nn.Conv2d) and isinstance(stem[1], nn.LayerNorm)):
        conv = _as_patch_conv(stem[0])
        if isinstance(conv, PatchConv2d):
            model.stem = StemB200(conv, _as_layernorm2d(stem[1]))
            n += 1
    for stage in getattr(model, "stages", []):
        ds = getattr(stage, "downsample", None)
        if (isinstance(ds, nn.Sequential) and not isinstance(ds, DownsampleB200) and len(ds) == 2
                and isinstance(ds[0], nn.LayerNorm) and isinstance(ds[1], nn.Conv2d)):
            conv = _as_patch_conv(ds[1])
            if isinstance(conv, PatchConv2d):
                stage.downsample = DownsampleB200(_as_layernorm2d(ds[0]), conv)
                n += 1
    head = getattr(model, "head", None)
    if head is not None:
        gp = getattr(head, "global_pool", None)
        # nn.AdaptiveAvgPool2d(1), or timm's SelectAdaptivePool2d('avg') that leaves flattening to
        # head.flatten
        is_avg1 = (isinstance(gp, nn.AdaptiveAvgPool2d) and gp.output_size in (1, (1, 1))) or \
                  (getattr(gp, "pool_type", None) == "avg"
                   and isinstance(getattr(gp, "flatten", None), nn.Identity))
        if is_avg1 and not isinstance(gp, GlobalAvgPoolB200):
            head.global_pool = GlobalAvgPoolB200(1)
            n += 1
        norm = getattr(head, "norm", None)
        if isinstance(norm, nn.LayerNorm) and not isinstance(norm, LayerNorm2d):
            head.norm = _as_layernorm2d(norm)
            n += 1
    return n


def create_convnext(name: str = "convnext_base", num_classes: int = 44, block_cls=ConvNeXtBlockB200,
                    ls_init_value: float = 1e-6) -> ConvNeXtTimmLike:
    """Stand-in for `timm.create_model(name, pretrained=False, num_classes=...)`."""
    cfg = CONVNEXT_CFGS[name]
    return ConvNeXtTimmLike(cfg["depths"], cfg["dims"], num_classes, block_cls=block_cls,
                            ls_init_value=ls_init_value)
