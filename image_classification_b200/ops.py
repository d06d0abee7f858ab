"""Tensor-level wrappers over the C ABI: argument checks, output allocation (by PyTorch's caching
allocator — the library owns no memory), current-stream plumbing.  Nothing here computes."""
from __future__ import annotations

import os
from dataclasses import dataclass

import torch

from . import _lib

LN_EPS = 1e-6  # timm ConvNeXtBlock norm eps


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _chk(t: torch.Tensor, dtype: torch.dtype, name: str) -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise _lib.CnxError(f"{name}: expected a CUDA tensor (libcnx_b200 has no CPU path)")
    if t.dtype != dtype:
        raise _lib.CnxError(f"{name}: expected dtype {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise _lib.CnxError(f"{name}: expected a contiguous tensor, got strides {t.stride()}")
    return t


def _p(t):
    return None if t is None else t.data_ptr()


# ------------------------------------------------------------------ weight folding
@dataclass
class FoldedWeights:
    """Output of cnx_prep_block_weights for one block (see csrc/prep.cu)."""

    wt_dw: torch.Tensor   # [49, C] fp32
    w1f: torch.Tensor     # [4C, C] bf16
    w1fT: torch.Tensor    # [C, 4C] bf16
    b1f: torch.Tensor     # [4C] fp32
    w2f: torch.Tensor     # [C, 4C] bf16
    w2fT: torch.Tensor    # [4C, C] bf16
    b2f: torch.Tensor     # [C] fp32
    rb: torch.Tensor      # [4C, 2] fp32: (rowsum of the bf16 w1f row, -b1f): LayerNorm backward by linearity


def prep_block_weights(conv_dw_w, norm_w, norm_b, fc1_w, fc1_b, fc2_w, fc2_b, gamma) -> FoldedWeights:
    C = gamma.numel()
    f32 = torch.float32
    for n, t in (("conv_dw.weight", conv_dw_w), ("norm.weight", norm_w), ("norm.bias", norm_b),
                 ("mlp.fc1.weight", fc1_w), ("mlp.fc1.bias", fc1_b), ("mlp.fc2.weight", fc2_w),
                 ("mlp.fc2.bias", fc2_b), ("gamma", gamma)):
        _chk(t, f32, n)
    dev = gamma.device
    fw = FoldedWeights(
        wt_dw=torch.empty(49, C, device=dev, dtype=f32),
        w1f=torch.empty(4 * C, C, device=dev, dtype=torch.bfloat16),
        w1fT=torch.empty(C, 4 * C, device=dev, dtype=torch.bfloat16),
        b1f=torch.empty(4 * C, device=dev, dtype=f32),
        w2f=torch.empty(C, 4 * C, device=dev, dtype=torch.bfloat16),
        w2fT=torch.empty(4 * C, C, device=dev, dtype=torch.bfloat16),
        b2f=torch.empty(C, device=dev, dtype=f32),
        rb=torch.empty(4 * C, 2, device=dev, dtype=f32),
    )
    _lib.call("cnx_prep_block_weights", _p(conv_dw_w), _p(norm_w), _p(norm_b), _p(fc1_w),
              _p(fc1_b), _p(fc2_w), _p(fc2_b), _p(gamma), _p(fw.wt_dw), _p(fw.w1f), _p(fw.w1fT),
              _p(fw.b1f), _p(fw.w2f), _p(fw.w2fT), _p(fw.b2f), _p(fw.rb), C, _stream(gamma))
    return fw


def unfold_block_grads(dw1f, db1f, dw2f, db2f, dwt_dw, norm_w, norm_b, fc1_w, fc2_w, fc2_b, gamma):
    """-> (d_conv_dw_w [C,1,7,7], d_norm_w, d_norm_b, d_fc1_w, d_fc2_w, d_fc2_b, d_gamma)."""
    C = gamma.numel()
    dev, f32 = gamma.device, torch.float32
    d_dw = torch.empty(C, 1, 7, 7, device=dev, dtype=f32)
    d_nw = torch.empty(C, device=dev, dtype=f32)
    d_nb = torch.empty(C, device=dev, dtype=f32)
    d_f1 = torch.empty(4 * C, C, device=dev, dtype=f32)
    d_f2 = torch.empty(C, 4 * C, device=dev, dtype=f32)
    d_f2b = torch.empty(C, device=dev, dtype=f32)
    d_g = torch.empty(C, device=dev, dtype=f32)
    _lib.call("cnx_unfold_block_grads", _p(dw1f), _p(db1f), _p(dw2f), _p(db2f), _p(dwt_dw),
              _p(norm_w), _p(norm_b), _p(fc1_w), _p(fc2_w), _p(fc2_b), _p(gamma), _p(d_dw),
              _p(d_nw), _p(d_nb), _p(d_f1), _p(d_f2), _p(d_f2b), _p(d_g), C, _stream(gamma))
    return d_dw, d_nw, d_nb, d_f1, d_f2, d_f2b, d_g


# ------------------------------------------------------------------ dwconv + LN
def dwconv_ln_fwd(x_nhwc: torch.Tensor, wt_dw: torch.Tensor, bias: torch.Tensor, eps: float = LN_EPS):
    """x [B,H,W,C] bf16 -> (xhat [B,H,W,C] bf16, mean [B*H*W] f32, rstd [B*H*W] f32)."""
    _chk(x_nhwc, torch.bfloat16, "x")
    _chk(wt_dw, torch.float32, "wt_dw")
    _chk(bias, torch.float32, "conv_dw.bias")
    B, H, W, C = x_nhwc.shape
    xhat = torch.empty_like(x_nhwc)
    mean = torch.empty(B * H * W, device=x_nhwc.device, dtype=torch.float32)
    rstd = torch.empty_like(mean)
    _lib.call("cnx_dwconv_ln_fwd", _p(x_nhwc), _p(wt_dw), _p(bias), _p(xhat), _p(mean), _p(rstd),
              B, H, W, C, eps, _stream(x_nhwc))
    if C <= 128 and not os.environ.get("CNX_DWCONV_NO_FUSE"):
        _lib.launch_count -= 1      # one fused kernel (LayerNorm in the conv epilogue), not two
    return xhat, mean, rstd


def dwconv_ln_bwd(dxhat, xhat, rstd, x_nhwc, dy, wt_dw, acc=None):
    """-> (dx [B,H,W,C] bf16, dwt [49,C] f32, dbias [C] f32, dy_colsum [C] f32).

    acc=(dwt, dbias, dy_colsum): accumulate (+=) into these existing fp32 buffers instead of
    overwriting fresh ones (gradient-accumulation window, see block.GradWindow)."""
    for n, t in (("dxhat", dxhat), ("xhat", xhat), ("x", x_nhwc), ("dy", dy)):
        _chk(t, torch.bfloat16, n)
    _chk(rstd, torch.float32, "rstd")
    _chk(wt_dw, torch.float32, "wt_dw")
    B, H, W, C = x_nhwc.shape
    dev = x_nhwc.device
    dx = torch.empty_like(x_nhwc)
    if acc is None:
        dwt = torch.empty(49, C, device=dev, dtype=torch.float32)
        dbias = torch.empty(C, device=dev, dtype=torch.float32)
        dycs = torch.empty(C, device=dev, dtype=torch.float32)
        entry = "cnx_dwconv_ln_bwd"
    else:
        dwt, dbias, dycs = (_chk(t, torch.float32, "acc") for t in acc)
        entry = "cnx_dwconv_ln_bwd_acc"
    # scratch for dconv = LN-backward(dxhat): sized by the library, owned by PyTorch's allocator
    ws_bytes = int(_lib.load().cnx_workspace_bytes(_lib.CNX_OP_DWCONV_LN_BWD, B, H, W, C))
    ws = torch.empty(ws_bytes, device=dev, dtype=torch.uint8)
    _lib.call(entry, _p(dxhat), _p(xhat), _p(rstd), _p(x_nhwc), _p(dy), _p(wt_dw),
              _p(dx), _p(dwt), _p(dbias), _p(dycs), B, H, W, C, _p(ws), ws_bytes, _stream(x_nhwc))
    return dx, dwt, dbias, dycs


# ------------------------------------------------------------------ LayerNorm2d (stem / downsample)
def layernorm_fwd(x2d, weight, bias, eps: float = LN_EPS):
    """x [M,C] bf16 -> (y [M,C] bf16, mean [M] f32, rstd [M] f32); affine LayerNorm over C."""
    _chk(x2d, torch.bfloat16, "x")
    _chk(weight, torch.float32, "weight")
    _chk(bias, torch.float32, "bias")
    M, C = x2d.shape
    y = torch.empty_like(x2d)
    mean = torch.empty(M, device=x2d.device, dtype=torch.float32)
    rstd = torch.empty_like(mean)
    _lib.call("cnx_layernorm_fwd", _p(x2d), _p(weight), _p(bias), _p(y), _p(mean), _p(rstd), M, C,
              eps, _stream(x2d))
    return y, mean, rstd


def layernorm_bwd(dy2d, x2d, weight, mean, rstd):
    """-> (dx [M,C] bf16, dweight [C] f32, dbias [C] f32)."""
    _chk(dy2d, torch.bfloat16, "dy")
    _chk(x2d, torch.bfloat16, "x")
    M, C = x2d.shape
    dx = torch.empty_like(x2d)
    dw = torch.empty(C, device=x2d.device, dtype=torch.float32)
    db = torch.empty_like(dw)
    _lib.call("cnx_layernorm_bwd", _p(dy2d), _p(x2d), _p(_chk(weight, torch.float32, "weight")),
              _p(mean), _p(rstd), _p(dx), _p(dw), _p(db), M, C, _stream(x2d))
    return dx, dw, db


def layernorm_patch_fwd(x_nhwc, weight, bias, patch: int, eps: float = LN_EPS):
    """LayerNorm2d of x [B,H,W,C] bf16 written patch-major: -> (y [B*H/k*W/k, k*k*C] bf16, mean, rstd)."""
    _chk(x_nhwc, torch.bfloat16, "x")
    B, H, W, C = x_nhwc.shape
    y = torch.empty(B * (H // patch) * (W // patch), patch * patch * C, device=x_nhwc.device,
                    dtype=torch.bfloat16)
    mean = torch.empty(B * H * W, device=x_nhwc.device, dtype=torch.float32)
    rstd = torch.empty_like(mean)
    _lib.call("cnx_layernorm_patch_fwd", _p(x_nhwc), _p(_chk(weight, torch.float32, "weight")),
              _p(_chk(bias, torch.float32, "bias")), _p(y), _p(mean), _p(rstd), B, H, W, C, patch,
              eps, _stream(x_nhwc))
    return y, mean, rstd


def layernorm_patch_bwd(dy_patch, x_nhwc, weight, mean, rstd, patch: int):
    """-> (dx [B,H,W,C] bf16, dweight [C], dbias [C]); dy_patch is patch-major like the forward's y."""
    _chk(dy_patch, torch.bfloat16, "dy")
    _chk(x_nhwc, torch.bfloat16, "x")
    B, H, W, C = x_nhwc.shape
    dx = torch.empty_like(x_nhwc)
    dw = torch.empty(C, device=x_nhwc.device, dtype=torch.float32)
    db = torch.empty_like(dw)
    _lib.call("cnx_layernorm_patch_bwd", _p(dy_patch), _p(x_nhwc), _p(_chk(weight, torch.float32, "weight")),
              _p(mean), _p(rstd), _p(dx), _p(dw), _p(db), B, H, W, C, patch, _stream(x_nhwc))
    return dx, dw, db


def patchify_nchw_f32(x: torch.Tensor, patch: int) -> torch.Tensor:
    """x [B,Cin,H,W] fp32 contiguous -> [B*H/k*W/k, k*k*Cin] bf16, inner order (kh, kw, c)."""
    _chk(x, torch.float32, "x")
    B, Cin, H, W = x.shape
    P = torch.empty(B * (H // patch) * (W // patch), patch * patch * Cin, device=x.device,
                    dtype=torch.bfloat16)
    _lib.call("cnx_patchify_nchw_f32", _p(x), _p(P), B, Cin, H, W, patch, _stream(x))
    return P


def dwconv_kxk_fwd(x: torch.Tensor, wt: torch.Tensor, k: int, stride: int) -> torch.Tensor:
    """x [B,H,W,C] bf16, wt [k*k, C] fp32 tap-major -> y [B,Ho,Wo,C] bf16 (padding k//2, no bias)."""
    _chk(x, torch.bfloat16, "x")
    _chk(wt, torch.float32, "wt")
    B, H, W, C = x.shape
    if wt.shape != (k * k, C):
        raise _lib.CnxError(f"dwconv_kxk_fwd: wt must be [{k * k}, {C}], got {tuple(wt.shape)}")
    p = k // 2
    Ho, Wo = (H + 2 * p - k) // stride + 1, (W + 2 * p - k) // stride + 1
    y = torch.empty(B, Ho, Wo, C, device=x.device, dtype=torch.bfloat16)
    _lib.call("cnx_dwconv_kxk_fwd", _p(x), _p(wt), _p(y), B, H, W, C, k, stride, _stream(x))
    return y


def dwconv_kxk_bwd(dy: torch.Tensor, x: torch.Tensor, wt: torch.Tensor, k: int, stride: int, *,
                   need_dx: bool = True, need_dw: bool = True, dwt_acc: torch.Tensor = None):
    """-> (dx [B,H,W,C] bf16 or None, dwt [k*k, C] fp32 or None).  dwt_acc: += into this buffer."""
    _chk(dy, torch.bfloat16, "dy")
    _chk(x, torch.bfloat16, "x")
    _chk(wt, torch.float32, "wt")
    B, H, W, C = x.shape
    p = k // 2
    Ho, Wo = (H + 2 * p - k) // stride + 1, (W + 2 * p - k) // stride + 1
    if dy.shape != (B, Ho, Wo, C) or wt.shape != (k * k, C):
        raise _lib.CnxError(f"dwconv_kxk_bwd: dy {tuple(dy.shape)} / wt {tuple(wt.shape)} do not "
                            f"match x {tuple(x.shape)}, k={k}, stride={stride}")
    dx = torch.empty_like(x) if need_dx else None
    dwt = None
    if need_dw:
        dwt = dwt_acc if dwt_acc is not None else torch.empty(k * k, C, device=x.device,
                                                              dtype=torch.float32)
        _chk(dwt, torch.float32, "dwt")
    _lib.call("cnx_dwconv_kxk_bwd", _p(dy), _p(x), _p(wt), _p(dx), _p(dwt), B, H, W, C, k, stride,
              int(dwt_acc is not None), _stream(x))
    return dx, dwt


def avgpool_nhwc_fwd(x: torch.Tensor) -> torch.Tensor:
    """[B, H, W, C] bf16 (contiguous NHWC) -> [B, C] bf16 mean over the pixels."""
    _chk(x, torch.bfloat16, "x")
    B, H, W, C = x.shape
    y = torch.empty(B, C, device=x.device, dtype=torch.bfloat16)
    _lib.call("cnx_avgpool_nhwc_fwd", _p(x), _p(y), B, H * W, C, _stream(x))
    return y


def avgpool_nhwc_bwd(g: torch.Tensor, H: int, W: int) -> torch.Tensor:
    """[B, C] bf16 gradient of the pooled map -> [B, H, W, C] bf16 (g / (H*W) at every pixel)."""
    _chk(g, torch.bfloat16, "g")
    B, C = g.shape
    dx = torch.empty(B, H, W, C, device=g.device, dtype=torch.bfloat16)
    _lib.call("cnx_avgpool_nhwc_bwd", _p(g), _p(dx), B, H * W, C, _stream(g))
    return dx


def colsum_bf16(x2d: torch.Tensor) -> torch.Tensor:
    """[M, C] bf16 -> [C] fp32 column sums."""
    _chk(x2d, torch.bfloat16, "x")
    M, C = x2d.shape
    out = torch.empty(C, device=x2d.device, dtype=torch.float32)
    _lib.call("cnx_colsum_bf16", _p(x2d), _p(out), M, C, _stream(x2d))
    return out


# ------------------------------------------------------------------ BatchNorm + SiLU, squeeze-excite
ACT_NONE, ACT_SILU = 0, 1


def bn_act_fwd(x: torch.Tensor, weight, bias, running_mean, running_var, *, training: bool,
               momentum: float, eps: float, act: int, want_pool: bool = False, out=None):
    """x [B,H,W,C] bf16 NHWC -> (y bf16 same shape, mean [C], rstd [C], pool_sum [B,C] fp32 | None).
    out: optional preallocated NHWC result (e.g. the permuted view of a channels_last tensor).
    training: batch statistics (cnx_bn_stats), running buffers updated in place when given;
    otherwise the running statistics normalise.  pool_sum = sum over the pixels of y per image."""
    _chk(x, torch.bfloat16, "x")
    B, H, W, C = x.shape
    f32 = torch.float32
    for n, t in (("weight", weight), ("bias", bias)):
        _chk(t, f32, n)
        if t.numel() != C:
            raise _lib.CnxError(f"bn_act_fwd: {n} must have {C} elements, got {t.numel()}")
    for n, t in (("running_mean", running_mean), ("running_var", running_var)):
        if t is not None:
            _chk(t, f32, n)
    dev = x.device
    sums = None
    if training:
        sums = torch.empty(2 * C, device=dev, dtype=f32)
        _lib.call("cnx_bn_stats", _p(x), _p(sums), B * H * W, C, _stream(x))
    if out is None:
        y = torch.empty_like(x)
    else:
        y = _chk(out, torch.bfloat16, "out")
        if y.shape != x.shape:
            raise _lib.CnxError(f"bn_act_fwd: out {tuple(y.shape)} vs x {tuple(x.shape)}")
    mean = torch.empty(C, device=dev, dtype=f32)
    rstd = torch.empty(C, device=dev, dtype=f32)
    pool = torch.empty(B, C, device=dev, dtype=f32) if want_pool else None
    _lib.call("cnx_bn_act_fwd", _p(x), _p(sums), _p(weight), _p(bias), _p(running_mean),
              _p(running_var), float(momentum), float(eps), int(act), _p(y), _p(mean), _p(rstd),
              _p(pool), B, H * W, C, _stream(x))
    return y, mean, rstd, pool


def bn_act_bwd(g: torch.Tensor, x: torch.Tensor, weight, bias, mean, rstd, act: int):
    """-> (dx bf16 like x, dweight [C] fp32, dbias [C] fp32) of y = act(bn(x)) with batch statistics."""
    _chk(g, torch.bfloat16, "g")
    _chk(x, torch.bfloat16, "x")
    if g.shape != x.shape:
        raise _lib.CnxError(f"bn_act_bwd: g {tuple(g.shape)} vs x {tuple(x.shape)}")
    C = x.shape[-1]
    M = x.numel() // C
    for n, t in (("weight", weight), ("bias", bias), ("mean", mean), ("rstd", rstd)):
        _chk(t, torch.float32, n)
    dx = torch.empty_like(x)
    dw = torch.empty(C, device=x.device, dtype=torch.float32)
    db = torch.empty(C, device=x.device, dtype=torch.float32)
    _lib.call("cnx_bn_act_bwd", _p(g), _p(x), _p(weight), _p(bias), _p(mean), _p(rstd), int(act),
              _p(dx), _p(dw), _p(db), M, C, _stream(x))
    return dx, dw, db


def se_scale_fwd(x: torch.Tensor, gate: torch.Tensor) -> torch.Tensor:
    """x [B,H,W,C] bf16, gate [B,C] fp32 -> x * gate[:, None, None, :] (bf16)."""
    _chk(x, torch.bfloat16, "x")
    _chk(gate, torch.float32, "gate")
    B, H, W, C = x.shape
    if tuple(gate.shape) != (B, C):
        raise _lib.CnxError(f"se_scale_fwd: gate must be [{B}, {C}], got {tuple(gate.shape)}")
    y = torch.empty_like(x)
    _lib.call("cnx_se_scale_fwd", _p(x), _p(gate), _p(y), B, H * W, C, _stream(x))
    return y


def se_scale_bwd_reduce(g: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    """-> dgate [B,C] fp32 = sum over pixels of g * x."""
    _chk(g, torch.bfloat16, "g")
    _chk(x, torch.bfloat16, "x")
    if g.shape != x.shape:
        raise _lib.CnxError(f"se_scale_bwd_reduce: g {tuple(g.shape)} vs x {tuple(x.shape)}")
    B, H, W, C = x.shape
    dgate = torch.empty(B, C, device=x.device, dtype=torch.float32)
    _lib.call("cnx_se_scale_bwd_reduce", _p(g), _p(x), _p(dgate), B, H * W, C, _stream(x))
    return dgate


def se_scale_bwd_apply(g: torch.Tensor, gate: torch.Tensor, dpool: torch.Tensor) -> torch.Tensor:
    """-> dx = g * gate + dpool / (H*W), bf16 like g (dpool [B,C] fp32: gradient at the pooled mean)."""
    _chk(g, torch.bfloat16, "g")
    _chk(gate, torch.float32, "gate")
    _chk(dpool, torch.float32, "dpool")
    B, H, W, C = g.shape
    if tuple(gate.shape) != (B, C) or tuple(dpool.shape) != (B, C):
        raise _lib.CnxError(f"se_scale_bwd_apply: gate / dpool must be [{B}, {C}]")
    dx = torch.empty_like(g)
    _lib.call("cnx_se_scale_bwd_apply", _p(g), _p(gate), _p(dpool), _p(dx), B, H * W, C, _stream(g))
    return dx


def se_gate_fwd(pool: torch.Tensor, pool_scale: float, w1, b1, w2t, b2):
    """pool [B,C] fp32 (pooled sums; pooled = pool * pool_scale), w1 [Cs,C], b1 [Cs], w2t [Cs,C]
    (expand weight transposed), b2 [C] -> (hidden [B,Cs] pre-activation, gate [B,C]), fp32."""
    for n, t in (("pool", pool), ("w1", w1), ("b1", b1), ("w2t", w2t), ("b2", b2)):
        _chk(t, torch.float32, n)
    B, C = pool.shape
    Cs = w1.shape[0]
    if tuple(w1.shape) != (Cs, C) or tuple(w2t.shape) != (Cs, C) or b1.numel() != Cs or b2.numel() != C:
        raise _lib.CnxError(f"se_gate_fwd: w1 / w2t must be [{Cs}, {C}], b1 [{Cs}], b2 [{C}]")
    hidden = torch.empty(B, Cs, device=pool.device, dtype=torch.float32)
    gate = torch.empty(B, C, device=pool.device, dtype=torch.float32)
    _lib.call("cnx_se_gate_fwd", _p(pool), float(pool_scale), _p(w1), _p(b1), _p(w2t), _p(b2),
              _p(hidden), _p(gate), B, C, Cs, _stream(pool))
    return hidden, gate


def se_gate_bwd(dgate, gate, hidden, pool, pool_scale: float, w1, w2t):
    """-> (dpool [B,C] gradient at the pooled vector, dw1 [Cs,C], db1 [Cs], dw2t [Cs,C], db2 [C])."""
    for n, t in (("dgate", dgate), ("gate", gate), ("hidden", hidden), ("pool", pool), ("w1", w1),
                 ("w2t", w2t)):
        _chk(t, torch.float32, n)
    B, C = pool.shape
    Cs = w1.shape[0]
    if tuple(dgate.shape) != (B, C) or tuple(gate.shape) != (B, C) or tuple(hidden.shape) != (B, Cs):
        raise _lib.CnxError(f"se_gate_bwd: dgate / gate must be [{B}, {C}], hidden [{B}, {Cs}]")
    dev, f32 = pool.device, torch.float32
    dz2 = torch.empty(B, C, device=dev, dtype=f32)
    dh = torch.empty(B, Cs, device=dev, dtype=f32)
    dpool = torch.empty(B, C, device=dev, dtype=f32)
    dw1 = torch.empty(Cs, C, device=dev, dtype=f32)
    db1 = torch.empty(Cs, device=dev, dtype=f32)
    dw2t = torch.empty(Cs, C, device=dev, dtype=f32)
    db2 = torch.empty(C, device=dev, dtype=f32)
    _lib.call("cnx_se_gate_bwd", _p(dgate), _p(gate), _p(hidden), _p(pool), float(pool_scale),
              _p(w1), _p(w2t), _p(dz2), _p(dh), _p(dpool), _p(dw1), _p(db1), _p(dw2t), _p(db2),
              B, C, Cs, _stream(pool))
    return dpool, dw1, db1, dw2t, db2


# ------------------------------------------------------------------ MLP GEMMs
EPI_STORE_BF16, EPI_BIAS_GELU, EPI_BIAS_RESIDUAL, EPI_GELU_GRAD, EPI_STORE_F32 = range(5)
EPI_BIAS = 6


def gemm_bf16(A, B, *, a_mn_major=False, b_mn_major=False, epilogue=EPI_STORE_BF16, bias=None,
              R=None, colsum=None):
    """D = epilogue(A @ B^T); A is [M,K] (or [K,M] if a_mn_major), B is [N,K] (or [K,N])."""
    _chk(A, torch.bfloat16, "A")
    _chk(B, torch.bfloat16, "B")
    if a_mn_major:
        K, M = A.shape
    else:
        M, K = A.shape
    if b_mn_major:
        Kb, N = B.shape
    else:
        N, Kb = B.shape
    if K != Kb:
        raise _lib.CnxError(f"gemm_bf16: K mismatch {K} vs {Kb}")
    dev = A.device
    out_dtype = torch.float32 if epilogue == EPI_STORE_F32 else torch.bfloat16
    if epilogue == EPI_STORE_F32:
        D0 = torch.zeros(M, N, device=dev, dtype=out_dtype)
    else:
        D0 = torch.empty(M, N, device=dev, dtype=out_dtype)
    D1 = torch.empty(M, N, device=dev, dtype=torch.bfloat16) if epilogue == EPI_BIAS_GELU else None
    if bias is not None:
        _chk(bias, torch.float32, "bias")
    if R is not None:
        _chk(R, torch.bfloat16, "R")
    if colsum is not None:
        _chk(colsum, torch.float32, "colsum")
    _lib.call("cnx_gemm_bf16", _p(A), A.shape[1], int(a_mn_major), _p(B), B.shape[1],
              int(b_mn_major), M, N, K, epilogue, _p(D0), N, _p(D1), N, _p(bias), _p(R),
              N if R is not None else 0, _p(colsum), _stream(A))
    return (D0, D1) if D1 is not None else D0


def fc1_gelu_fwd(xhat2d, w1f, b1f):
    M, C = xhat2d.shape
    z = torch.empty(M, 4 * C, device=xhat2d.device, dtype=torch.bfloat16)
    h = torch.empty_like(z)
    _lib.call("cnx_fc1_gelu_fwd", _p(_chk(xhat2d, torch.bfloat16, "xhat")),
              _p(_chk(w1f, torch.bfloat16, "w1f")), _p(_chk(b1f, torch.float32, "b1f")), _p(z),
              _p(h), M, C, _stream(xhat2d))
    return z, h


def fc2_residual_fwd(h, w2f, b2f, x2d):
    M, C = x2d.shape
    y = torch.empty_like(x2d)
    _lib.call("cnx_fc2_residual_fwd", _p(_chk(h, torch.bfloat16, "h")),
              _p(_chk(w2f, torch.bfloat16, "w2f")), _p(_chk(b2f, torch.float32, "b2f")),
              _p(_chk(x2d, torch.bfloat16, "x")), _p(y), M, C, _stream(x2d))
    return y


def mlp_bwd_dz(dy2d, w2fT, z, db1f=None):
    """db1f: existing fp32 [4C] accumulator (+=) or None for a fresh zero-filled one."""
    M, C = dy2d.shape
    dz = torch.empty_like(z)
    if db1f is None:
        db1f = torch.zeros(4 * C, device=z.device, dtype=torch.float32)
    _lib.call("cnx_mlp_bwd_dz", _p(_chk(dy2d, torch.bfloat16, "dy")),
              _p(_chk(w2fT, torch.bfloat16, "w2fT")), _p(_chk(z, torch.bfloat16, "z")), _p(dz),
              _p(db1f), M, C, _stream(z))
    return dz, db1f


def mlp_bwd_dz_stats(dy2d, w2fT, z, rb, db1f=None):
    """cnx_mlp_bwd_dz that also returns rowstats [M,2] fp32 = (sum_c dxhat, sum_c dxhat * xhat)
    per token, taken in the GEMM epilogue (see include/cnx_b200.h).  -> (dz, db1f, rowstats)."""
    M, C = dy2d.shape
    dz = torch.empty_like(z)
    if db1f is None:
        db1f = torch.zeros(4 * C, device=z.device, dtype=torch.float32)
    rowstats = torch.zeros(M, 2, device=z.device, dtype=torch.float32)
    _lib.call("cnx_mlp_bwd_dz_stats", _p(_chk(dy2d, torch.bfloat16, "dy")),
              _p(_chk(w2fT, torch.bfloat16, "w2fT")), _p(_chk(z, torch.bfloat16, "z")),
              _p(_chk(rb, torch.float32, "rb")), _p(dz), _p(db1f), _p(rowstats), M, C, _stream(z))
    return dz, db1f, rowstats


def mlp_bwd_dconv(dz, w1fT, xhat2d, rowstats, rstd):
    """fc1 data-gradient GEMM with the LayerNorm backward in its epilogue -> dconv [M,C] bf16."""
    M, C4 = dz.shape
    C = C4 // 4
    dconv = torch.empty(M, C, device=dz.device, dtype=torch.bfloat16)
    _lib.call("cnx_mlp_bwd_dconv", _p(_chk(dz, torch.bfloat16, "dz")),
              _p(_chk(w1fT, torch.bfloat16, "w1fT")), _p(_chk(xhat2d, torch.bfloat16, "xhat")),
              _p(_chk(rowstats, torch.float32, "rowstats")), _p(_chk(rstd, torch.float32, "rstd")),
              _p(dconv), M, C, _stream(dz))
    return dconv


def dwconv_bwd(dconv, x_nhwc, dy, wt_dw, acc=None):
    """Depthwise 7x7 conv backward alone -> (dx, dwt [49,C], dbias [C], dy_colsum [C]);
    acc=(dwt, dbias, dy_colsum): += into these buffers."""
    for n, t in (("dconv", dconv), ("x", x_nhwc), ("dy", dy)):
        _chk(t, torch.bfloat16, n)
    _chk(wt_dw, torch.float32, "wt_dw")
    B, H, W, C = x_nhwc.shape
    dev = x_nhwc.device
    dx = torch.empty_like(x_nhwc)
    if acc is None:
        dwt = torch.empty(49, C, device=dev, dtype=torch.float32)
        dbias = torch.empty(C, device=dev, dtype=torch.float32)
        dycs = torch.empty(C, device=dev, dtype=torch.float32)
    else:
        dwt, dbias, dycs = (_chk(t, torch.float32, "acc") for t in acc)
    _lib.call("cnx_dwconv_bwd", _p(dconv), _p(x_nhwc), _p(dy), _p(wt_dw), _p(dx), _p(dwt), _p(dbias),
              _p(dycs), B, H, W, C, int(acc is not None), _stream(x_nhwc))
    return dx, dwt, dbias, dycs


def mlp_bwd_dxhat(dz, w1fT):
    M, C4 = dz.shape
    C = C4 // 4
    dxhat = torch.empty(M, C, device=dz.device, dtype=torch.bfloat16)
    _lib.call("cnx_mlp_bwd_dxhat", _p(_chk(dz, torch.bfloat16, "dz")),
              _p(_chk(w1fT, torch.bfloat16, "w1fT")), _p(dxhat), M, C, _stream(dz))
    return dxhat


def mlp_bwd_wgrad(dy2d, h, dz, xhat2d, acc=None):
    """acc=(dw2f, dw1f): existing fp32 accumulators (+=) or None for fresh zero-filled ones."""
    M, C = dy2d.shape
    if acc is None:
        dw2f = torch.zeros(C, 4 * C, device=h.device, dtype=torch.float32)
        dw1f = torch.zeros(4 * C, C, device=h.device, dtype=torch.float32)
    else:
        dw2f, dw1f = (_chk(t, torch.float32, "acc") for t in acc)
    _lib.call("cnx_mlp_bwd_wgrad", _p(_chk(dy2d, torch.bfloat16, "dy")),
              _p(_chk(h, torch.bfloat16, "h")), _p(_chk(dz, torch.bfloat16, "dz")),
              _p(_chk(xhat2d, torch.bfloat16, "xhat")), _p(dw2f), _p(dw1f), M, C, _stream(h))
    return dw2f, dw1f


# ------------------------------------------------------------------ layout
def nchw_f32_to_nhwc_bf16(x: torch.Tensor) -> torch.Tensor:
    _chk(x, torch.float32, "x")
    B, C, H, W = x.shape
    out = torch.empty(B, H, W, C, device=x.device, dtype=torch.bfloat16)
    _lib.call("cnx_nchw_f32_to_nhwc_bf16", _p(x), _p(out), B, C, H, W, _stream(x))
    return out


def nhwc_bf16_to_nchw_f32(x: torch.Tensor) -> torch.Tensor:
    _chk(x, torch.bfloat16, "x")
    B, H, W, C = x.shape
    out = torch.empty(B, C, H, W, device=x.device, dtype=torch.float32)
    _lib.call("cnx_nhwc_bf16_to_nchw_f32", _p(x), _p(out), B, C, H, W, _stream(x))
    return out
