"""ConvNeXtBlockB200 — the drop-in `torch.nn.Module` for timm's `ConvNeXtBlock`.

Boundary: same constructor meaning, same parameter/submodule names as timm
(`conv_dw.{weight,bias}` with `.out_channels` — read at /root/reference/train_advanced_v4.py:108-109,
`norm.{weight,bias}`, `mlp.fc1.{weight,bias}`, `mlp.fc2.{weight,bias}`, `gamma`), NCHW in / NCHW
out, fp32 master parameters and fp32 parameter gradients, so `state_dict()`, `EMA`, AdamW,
`clip_grad_norm_` and checkpoints (/root/reference/train_advanced_v4.py:71,228,466,506-510) see
nothing new.  All arithmetic happens in libcnx_b200 (five kernels forward incl. weight folding,
six backward); this file only routes tensors.
"""
from __future__ import annotations

import os

import torch
from torch import nn

from . import _lib, ops

# Backward: the two weight-gradient GEMMs of a block depend only on tensors that exist once the dz
# GEMM has run, and nothing downstream in the block needs their results — so they are enqueued on a
# second stream and the data-gradient chain (dxhat GEMM -> LayerNorm-bwd -> conv-bwd) continues on
# the main one; the streams join at the end of the block's backward.  Persistent kernels fill each
# other's tails and the memory-bound row kernels run under the tensor-bound GEMMs.
# CNX_BWD_SIDE_STREAM=0 keeps everything on one stream (A/B switch).
_side_streams: dict = {}


def _side_stream(device: torch.device):
    if os.environ.get("CNX_BWD_SIDE_STREAM", "1") == "0":
        return None
    key = device.index if device.index is not None else torch.cuda.current_device()
    if key not in _side_streams:
        _side_streams[key] = torch.cuda.Stream(device)
    return _side_streams[key]


# Bumped by every writer that changes parameter VALUES behind autograd's back (the fused optimizer
# step updates parameters from a ctypes-launched kernel through raw pointers, so `p._version` does
# not move): part of the folded-weight cache key of every block.
_weight_epoch = 0


# While a CUDA graph is being captured the fold is always recorded (a replay then folds the weights
# of THAT replay).  A step made of several graphs replayed in a fixed order (one per micro-batch of
# an accumulation window, graphed.GraphedTrainStep) records it in the first graph only and sets
# this to True while capturing the others, which then read the tensors the first graph writes.
capture_reuses_fold = False


def notify_weights_changed() -> None:
    """Invalidate the folded-weight caches of all ConvNeXtBlockB200 modules.  Call after writing
    parameter storage with anything that is not a torch in-place op (raw-pointer kernels)."""
    global _weight_epoch
    _weight_epoch += 1


_fold_event = None          # recorded on the side stream by prefold_blocks(); consumed by the first block


def prefold_blocks(blocks, device) -> None:
    """Fold the weights of every block in `blocks` NOW, on the side stream, instead of one by one
    as the forward pass reaches them: the ~36 small fold kernels then run under the stem (which
    needs none of them).  The first block forward waits for the side stream.  Called by
    `train_step.TrainStep` at the start of every accumulation window."""
    global _fold_event, capture_reuses_fold
    side = _side_stream(device)
    if side is None or not blocks:
        return
    main = torch.cuda.current_stream(device)
    side.wait_stream(main)
    with torch.cuda.stream(side):
        for b in blocks:
            b._folded(b.conv_dw.weight, b.norm.weight, b.norm.bias, b.mlp.fc1.weight,
                      b.mlp.fc1.bias, b.mlp.fc2.weight, b.mlp.fc2.bias, b.gamma)
        ev = torch.cuda.Event()
        ev.record(side)
    _fold_event = ev
    if torch.cuda.is_current_stream_capturing():
        capture_reuses_fold = True      # the rest of this capture reads the tensors just recorded


def _l2_sub_batches(B: int, H: int, W: int, C: int) -> int:
    """L2 blocking over the batch — an experiment kept behind CNX_L2_CHUNK_MB (default 0 = off).

    At the first two stages one `[M, 4C]` hidden tensor of a 128-image micro-batch (411 / 205 MB) is
    several times the 126 MB L2, so every tensor a kernel of the block hands to the next one makes a
    round trip through HBM.  Nothing in the block mixes samples (LayerNorm is per token), so the
    block CAN run sub-batch by sub-batch through all its kernels, sized so that a sub-batch's hidden
    tensor and what travels with it stay L2-resident between producer and consumer; in backward
    `dz`, `dxhat` and `dconv` then live in chunk-sized buffers that are overwritten while still
    dirty in L2.  Measured (same box, `tools/gpu_chunk_ab.sh`, profiles/r02_l2_chunk_ab.txt): with
    ~50 MB chunks (8 sub-batches at stage 1, 4 at stage 2: +558 launches per step) the step takes
    46.1 ms against 42.8 ms unchunked, with 28 MB chunks 49.7 ms — every launch costs its ~6 us of
    prologue / tail whatever its size and the smaller GEMMs lose more to wave quantisation than the
    L2 hits give back.  Off by default.  Returns the number of sub-batches (a divisor of B)."""
    mb = float(os.environ.get("CNX_L2_CHUNK_MB", "0"))
    hidden_mb = B * H * W * 4 * C * 2 / 1e6
    if mb <= 0 or C > 256 or hidden_mb <= 2.0 * mb:
        return 1
    for d in range(2, B + 1):
        if B % d == 0 and hidden_mb / d <= mb:
            return d
    return 1


class _Mlp(nn.Module):
    """Parameter container named like timm's `Mlp` (fc1 / act / fc2)."""

    def __init__(self, dim: int, hidden: int):
        super().__init__()
        self.fc1 = nn.Linear(dim, hidden)
        self.act = nn.GELU()
        self.fc2 = nn.Linear(hidden, dim)


def _forward_sub_batched(x_nhwc, fw, conv_b, nsub):
    """The three forward kernels of a block, sub-batch by sub-batch (see _l2_sub_batches); full-size
    outputs, every call works on a batch slice (contiguous in NHWC)."""
    B, H, W, C = x_nhwc.shape
    M, Bs = B * H * W, B // nsub
    Ms = Bs * H * W
    dev, bf16 = x_nhwc.device, torch.bfloat16
    xhat = torch.empty_like(x_nhwc)
    mean = torch.empty(M, device=dev, dtype=torch.float32)
    rstd = torch.empty_like(mean)
    z = torch.empty(M, 4 * C, device=dev, dtype=bf16)
    h = torch.empty_like(z)
    y = torch.empty(M, C, device=dev, dtype=bf16)
    x2d, xh2d = x_nhwc.view(M, C), xhat.view(M, C)
    st = torch.cuda.current_stream(dev).cuda_stream
    fused = C <= 128 and not os.environ.get("CNX_DWCONV_NO_FUSE")
    for i in range(nsub):
        r0, r1 = i * Ms, (i + 1) * Ms
        _lib.call("cnx_dwconv_ln_fwd", x2d[r0:r1].data_ptr(), fw.wt_dw.data_ptr(), conv_b.data_ptr(),
                  xh2d[r0:r1].data_ptr(), mean[r0:r1].data_ptr(), rstd[r0:r1].data_ptr(),
                  Bs, H, W, C, ops.LN_EPS, st)
        if fused:
            _lib.launch_count -= 1
        _lib.call("cnx_fc1_gelu_fwd", xh2d[r0:r1].data_ptr(), fw.w1f.data_ptr(), fw.b1f.data_ptr(),
                  z[r0:r1].data_ptr(), h[r0:r1].data_ptr(), Ms, C, st)
        _lib.call("cnx_fc2_residual_fwd", h[r0:r1].data_ptr(), fw.w2f.data_ptr(), fw.b2f.data_ptr(),
                  x2d[r0:r1].data_ptr(), y[r0:r1].data_ptr(), Ms, C, st)
    return xhat, rstd, z, h, y


def _backward_sub_batched(x_nhwc, xhat, rstd, z, h, dy_nhwc, fw, a, nsub):
    """Backward of a block sub-batch by sub-batch: dz, dxhat and the LayerNorm-backward scratch live
    in chunk-sized buffers that are rewritten while still dirty in L2 (they never reach HBM);
    every parameter-gradient output accumulates (+=) over the sub-batches."""
    B, H, W, C = x_nhwc.shape
    M, Bs = B * H * W, B // nsub
    Ms = Bs * H * W
    dev, bf16, f32 = x_nhwc.device, torch.bfloat16, torch.float32
    if a is None:
        dw2f = torch.zeros(C, 4 * C, device=dev, dtype=f32)
        dw1f = torch.zeros(4 * C, C, device=dev, dtype=f32)
        db1f = torch.zeros(4 * C, device=dev, dtype=f32)
        dwt = torch.zeros(49, C, device=dev, dtype=f32)
        dbias = torch.zeros(C, device=dev, dtype=f32)
        db2f = torch.zeros(C, device=dev, dtype=f32)
    else:
        dw2f, dw1f, db1f, dwt, dbias, db2f = (a["dw2f"], a["dw1f"], a["db1f"], a["dwt"], a["dbias"],
                                              a["db2f"])
    dx = torch.empty_like(x_nhwc)
    dz_c = torch.empty(Ms, 4 * C, device=dev, dtype=bf16)
    dxhat_c = torch.empty(Ms, C, device=dev, dtype=bf16)
    ws_bytes = int(_lib.load().cnx_workspace_bytes(_lib.CNX_OP_DWCONV_LN_BWD, Bs, H, W, C))
    ws = torch.empty(ws_bytes, device=dev, dtype=torch.uint8)
    x2d, xh2d, dy2d, dx2d = x_nhwc.view(M, C), xhat.view(M, C), dy_nhwc.view(M, C), dx.view(M, C)
    st = torch.cuda.current_stream(dev).cuda_stream
    for i in range(nsub):
        r0, r1 = i * Ms, (i + 1) * Ms
        _lib.call("cnx_mlp_bwd_dz", dy2d[r0:r1].data_ptr(), fw.w2fT.data_ptr(), z[r0:r1].data_ptr(),
                  dz_c.data_ptr(), db1f.data_ptr(), Ms, C, st)
        _lib.call("cnx_mlp_bwd_dxhat", dz_c.data_ptr(), fw.w1fT.data_ptr(), dxhat_c.data_ptr(), Ms, C, st)
        _lib.call("cnx_mlp_bwd_wgrad", dy2d[r0:r1].data_ptr(), h[r0:r1].data_ptr(), dz_c.data_ptr(),
                  xh2d[r0:r1].data_ptr(), dw2f.data_ptr(), dw1f.data_ptr(), Ms, C, st)
        _lib.call("cnx_dwconv_ln_bwd_acc", dxhat_c.data_ptr(), xh2d[r0:r1].data_ptr(),
                  rstd[r0:r1].data_ptr(), x2d[r0:r1].data_ptr(), dy2d[r0:r1].data_ptr(),
                  fw.wt_dw.data_ptr(), dx2d[r0:r1].data_ptr(), dwt.data_ptr(), dbias.data_ptr(),
                  db2f.data_ptr(), Bs, H, W, C, ws.data_ptr(), ws_bytes, st)
    return dx, dw2f, dw1f, db1f, dwt, dbias, db2f


class _BlockFn(torch.autograd.Function):
    """x: [B,C,H,W] bf16, channels_last (physically NHWC).  Returns the same layout."""

    @staticmethod
    def forward(ctx, x, conv_w, conv_b, norm_w, norm_b, fc1_w, fc1_b, fc2_w, fc2_b, gamma, module):
        B, C, H, W = x.shape
        x_nhwc = x.permute(0, 2, 3, 1)
        assert x_nhwc.is_contiguous() and x.dtype == torch.bfloat16
        global _fold_event
        if _fold_event is not None:          # folds enqueued on the side stream (prefold_blocks)
            torch.cuda.current_stream(x.device).wait_event(_fold_event)
            _fold_event = None
        fw = module._folded(conv_w, norm_w, norm_b, fc1_w, fc1_b, fc2_w, fc2_b, gamma)
        M = B * H * W
        nsub = _l2_sub_batches(B, H, W, C)
        if nsub == 1:
            xhat, _mean, rstd = ops.dwconv_ln_fwd(x_nhwc, fw.wt_dw, conv_b.detach())
            z, h = ops.fc1_gelu_fwd(xhat.view(M, C), fw.w1f, fw.b1f)
            y = ops.fc2_residual_fwd(h, fw.w2f, fw.b2f, x_nhwc.reshape(M, C))
        else:
            xhat, rstd, z, h, y = _forward_sub_batched(x_nhwc, fw, conv_b.detach(), nsub)
        ctx.nsub = nsub
        ctx.save_for_backward(x, xhat, rstd, z, h, norm_w, norm_b, fc1_w, fc2_w, fc2_b, gamma)
        ctx.fw = fw
        ctx.module = module
        return y.view(B, H, W, C).permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        x, xhat, rstd, z, h, norm_w, norm_b, fc1_w, fc2_w, fc2_b, gamma = ctx.saved_tensors
        fw = ctx.fw
        B, C, H, W = x.shape
        M = B * H * W
        dy_nhwc = dy.permute(0, 2, 3, 1)
        if dy_nhwc.dtype != torch.bfloat16 or not dy_nhwc.is_contiguous():
            dy_nhwc = dy_nhwc.to(torch.bfloat16).contiguous()
        dy2d = dy_nhwc.view(M, C)
        win = ctx.module._window          # GradWindow slot of this block, or None
        a = win.acc if win is not None and win.owner.active else None
        if ctx.nsub > 1:
            dx, dw2f, dw1f, db1f, dwt, dbias, db2f = _backward_sub_batched(
                x.permute(0, 2, 3, 1), xhat, rstd, z, h, dy_nhwc, fw, a, ctx.nsub)
        else:
            # CNX_LNBWD_LINEARITY=1: LayerNorm backward by linearity (include/cnx_b200.h,
            # cnx_mlp_bwd_dz_stats): the per-token sums it needs come out of the dz GEMM's epilogue,
            # the fc1 data-gradient GEMM applies them in its own epilogue and hands the conv its
            # output gradient directly — no LayerNorm-backward row kernel, no dxhat / dconv round trip
            # (stage-1 DRAM traffic of the LayerNorm part 516 -> 309 MB, 144 launches fewer per step,
            # half the rounding error).  Measured per stage (tools/lnbwd_microbench.py,
            # profiles/r02_lnbwd_linearity_ab.txt): 637 / 363 / 232 / 194 us against 629 / 356 / 228 /
            # 200 us for dz GEMM + dxhat GEMM + row kernel + conv-bwd — the three extra packed FMAs per
            # two elements in the FMA-pipe-bound dz epilogue cost what the row kernel cost, and the
            # step time is unchanged (43.7-44.2 vs 43.4-43.8 ms), so the default stays the row kernel.
            rowk = os.environ.get("CNX_LNBWD_LINEARITY") != "1"
            if rowk:
                dz, db1f = ops.mlp_bwd_dz(dy2d, fw.w2fT, z, None if a is None else a["db1f"])
            else:
                dz, db1f, rowstats = ops.mlp_bwd_dz_stats(dy2d, fw.w2fT, z, fw.rb,
                                                          None if a is None else a["db1f"])
            side = _side_stream(x.device)
            if side is None:
                dw2f, dw1f = ops.mlp_bwd_wgrad(dy2d, h, dz, xhat.view(M, C),
                                               None if a is None else (a["dw2f"], a["dw1f"]))
            else:
                # weight gradients on the side stream; their outputs are allocated here, on the main one
                if a is None:
                    wacc = (torch.zeros(C, 4 * C, device=x.device, dtype=torch.float32),
                            torch.zeros(4 * C, C, device=x.device, dtype=torch.float32))
                else:
                    wacc = (a["dw2f"], a["dw1f"])
                main = torch.cuda.current_stream(x.device)
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    dw2f, dw1f = ops.mlp_bwd_wgrad(dy2d, h, dz, xhat.view(M, C), wacc)
            cacc = None if a is None else (a["dwt"], a["dbias"], a["db2f"])
            if rowk:
                dxhat = ops.mlp_bwd_dxhat(dz, fw.w1fT)
                dx, dwt, dbias, db2f = ops.dwconv_ln_bwd(dxhat.view(B, H, W, C), xhat, rstd,
                                                         x.permute(0, 2, 3, 1), dy_nhwc, fw.wt_dw, cacc)
            else:
                dconv = ops.mlp_bwd_dconv(dz, fw.w1fT, xhat.view(M, C), rowstats, rstd)
                dx, dwt, dbias, db2f = ops.dwconv_bwd(dconv.view(B, H, W, C), x.permute(0, 2, 3, 1),
                                                      dy_nhwc, fw.wt_dw, cacc)
            if side is not None:
                torch.cuda.current_stream(x.device).wait_stream(side)
        if a is not None and not win.owner.boundary:
            # inside an accumulation window: the folded-parameter gradients stay in the window's
            # fp32 accumulators; parameter gradients are produced on the boundary micro-batch only
            return (dx.permute(0, 3, 1, 2),) + (None,) * 10
        d_conv_w, d_norm_w, d_norm_b, d_fc1_w, d_fc2_w, d_fc2_b, d_gamma = ops.unfold_block_grads(
            dw1f, db1f, dw2f, db2f, dwt, norm_w.detach(), norm_b.detach(), fc1_w.detach(),
            fc2_w.detach(), fc2_b.detach(), gamma.detach())
        if a is not None:
            # the two pass-through gradients alias the window's accumulators: hand out copies
            dbias, db1f = dbias.clone(), db1f.clone()
        return (dx.permute(0, 3, 1, 2), d_conv_w, dbias, d_norm_w, d_norm_b, d_fc1_w, db1f,
                d_fc2_w, d_fc2_b, d_gamma, None)


class _NchwF32ToNhwcBf16(torch.autograd.Function):
    """fp32 contiguous NCHW -> bf16 channels_last, and the reverse for the gradient."""

    @staticmethod
    def forward(ctx, x):
        return ops.nchw_f32_to_nhwc_bf16(x).permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, g):
        g = g.permute(0, 2, 3, 1)
        if g.dtype != torch.bfloat16 or not g.is_contiguous():
            g = g.to(torch.bfloat16).contiguous()
        return ops.nhwc_bf16_to_nchw_f32(g)


class _NhwcBf16ToNchwF32(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y):
        return ops.nhwc_bf16_to_nchw_f32(y.permute(0, 2, 3, 1))

    @staticmethod
    def backward(ctx, g):
        return ops.nchw_f32_to_nhwc_bf16(g.float().contiguous()).permute(0, 3, 1, 2)


class ConvNeXtBlockB200(nn.Module):
    """dwconv7x7 -> LayerNorm(C) -> Linear C->4C -> GELU -> Linear 4C->C -> *gamma -> +shortcut.

    `forward` accepts what a timm stage hands it:
      * bf16 channels_last  -> zero-copy, returns bf16 channels_last (the fast path a
        channels_last bf16 model such as `convnext.ConvNeXtTimmLike` uses end to end);
      * fp32 contiguous NCHW (the reference's autocast residual stream) -> converted by the
        library's own transpose kernels on the way in and out, returns fp32 NCHW;
      * anything else       -> converted with torch ops to bf16 channels_last and back.
    """

    def __init__(self, dim: int, ls_init_value: float | None = 1e-6, mlp_ratio: int = 4):
        super().__init__()
        if mlp_ratio != 4:
            raise ValueError("ConvNeXtBlockB200 implements the mlp_ratio=4 block only")
        self.conv_dw = nn.Conv2d(dim, dim, kernel_size=7, padding=3, groups=dim, bias=True)
        self.norm = nn.LayerNorm(dim, eps=ops.LN_EPS)
        self.mlp = _Mlp(dim, 4 * dim)
        init = 1.0 if ls_init_value is None else ls_init_value
        self.gamma = nn.Parameter(init * torch.ones(dim), requires_grad=ls_init_value is not None)
        self.drop_path = nn.Identity()   # create_model passes no drop_path_rate (:488-492)
        self._fold_key = None
        self._fold = None
        self._window = None              # set by GradWindow

    # -- folded bf16 weights, cached per weight version -------------------------------------
    def _folded(self, *params) -> ops.FoldedWeights:
        # key: storage + torch's version counter (in-place torch ops, optimizer.step of torch's own
        # optimizers) + the global epoch raw-pointer writers bump (optim_step.FusedOptimStep).
        # Inside a CUDA-graph capture the fold is always recorded, so a replay refolds the weights
        # of that replay instead of baking a stale copy into the graph.
        key = (_weight_epoch,) + tuple((p.data_ptr(), p._version) for p in params)
        capturing = torch.cuda.is_current_stream_capturing()
        if capturing and capture_reuses_fold and self._fold is not None:
            return self._fold
        if self._fold is None or key != self._fold_key or capturing:
            self._fold = ops.prep_block_weights(*[p.detach() for p in params])
            self._fold_key = key
        return self._fold

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not x.is_cuda:
            raise _lib.CnxError("ConvNeXtBlockB200 runs on sm_100a CUDA devices only; use the "
                                "oracle block for CPU reference computations")
        mode = 0
        if x.dtype == torch.bfloat16 and x.permute(0, 2, 3, 1).is_contiguous():
            xin = x
        elif x.dtype == torch.float32 and x.is_contiguous():
            mode = 1
            xin = _NchwF32ToNhwcBf16.apply(x)
        else:
            mode = 2
            xin = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
            if not xin.permute(0, 2, 3, 1).is_contiguous():   # degenerate 1x1 / C==1 stride cases
                xin = xin.permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)
        y = _BlockFn.apply(xin, self.conv_dw.weight, self.conv_dw.bias, self.norm.weight,
                           self.norm.bias, self.mlp.fc1.weight, self.mlp.fc1.bias,
                           self.mlp.fc2.weight, self.mlp.fc2.bias, self.gamma, self)
        if mode == 1:
            return _NhwcBf16ToNchwF32.apply(y)
        if mode == 2:
            return y.to(x.dtype)
        return y

    @torch.no_grad()
    def load_from(self, other: nn.Module) -> "ConvNeXtBlockB200":
        """Copy parameters from any module exposing timm's ConvNeXtBlock attribute names."""
        self.conv_dw.weight.copy_(other.conv_dw.weight)
        self.conv_dw.bias.copy_(other.conv_dw.bias)
        self.norm.weight.copy_(other.norm.weight)
        self.norm.bias.copy_(other.norm.bias)
        self.mlp.fc1.weight.copy_(other.mlp.fc1.weight)
        self.mlp.fc1.bias.copy_(other.mlp.fc1.bias)
        self.mlp.fc2.weight.copy_(other.mlp.fc2.weight)
        self.mlp.fc2.bias.copy_(other.mlp.fc2.bias)
        if getattr(other, "gamma", None) is not None:
            self.gamma.copy_(other.gamma.reshape(-1))
        return self


class _WindowSlot:
    __slots__ = ("owner", "acc")

    def __init__(self, owner, acc):
        self.owner, self.acc = owner, acc


class GradWindow:
    """Gradient-accumulation window for every ConvNeXtBlockB200 of `model`.

    The reference accumulates 2 micro-batches per optimizer step
    (/root/reference/train_advanced_v4.py:225, config.py:37).  Autograd would run the adjoint of the
    weight folding (cnx_unfold_block_grads) and an `AccumulateGrad` add per parameter on every
    micro-batch.  Inside a window the blocks instead accumulate the gradients of their FOLDED
    parameters (fp32, one flat buffer for the whole model, zeroed by one memset) across
    micro-batches through the kernels' += outputs, and emit parameter gradients only on the
    boundary micro-batch.  The sum is the same; only the summation order differs.

        win = GradWindow(model)
        win.begin(); win.set_boundary(False); loss1.backward()
        win.set_boundary(True); loss2.backward()      # .grad of the block parameters appears here
    """

    def __init__(self, model: nn.Module):
        self.blocks = [m for m in model.modules() if isinstance(m, ConvNeXtBlockB200)]
        self.active = False
        self.boundary = True
        self.flat = None
        if not self.blocks:
            return
        dev = self.blocks[0].gamma.device
        shapes = lambda C: (("dw1f", (4 * C, C)), ("dw2f", (C, 4 * C)), ("db1f", (4 * C,)),
                            ("dwt", (49, C)), ("dbias", (C,)), ("db2f", (C,)))
        total = 0
        for b in self.blocks:
            C = b.gamma.numel()
            total += sum((_numel(sh) + 3) // 4 * 4 for _, sh in shapes(C))
        self.flat = torch.zeros(total, device=dev, dtype=torch.float32)
        off = 0
        for b in self.blocks:
            acc = {}
            for name, sh in shapes(b.gamma.numel()):
                n = _numel(sh)
                acc[name] = self.flat[off:off + n].view(sh)
                off += (n + 3) // 4 * 4          # keep every slice 16-byte aligned
            b._window = _WindowSlot(self, acc)

    def begin(self):
        """Start a window: zero all accumulators (one memset).  A window that never reached its
        boundary (an odd trailing micro-batch: the reference leaves those gradients in `.grad` and
        they join the next optimizer step, /root/reference/train_advanced_v4.py:225-231) is
        continued, not zeroed."""
        if self.flat is not None:
            if not self.active:
                self.flat.zero_()
            self.active = True
            self.boundary = False

    def set_boundary(self, boundary: bool):
        self.boundary = bool(boundary)

    def end(self):
        """Close the window (called after the boundary micro-batch): later backward passes outside
        a window take the plain per-call path instead of adding stale accumulators."""
        self.active = False
        self.boundary = True

    @torch.no_grad()
    def flush(self):
        """Emit what an unfinished window holds into the parameters' `.grad` (adding to what is
        there), exactly where the reference's plain accumulation would have left it, and close."""
        if self.flat is None or not self.active:
            return
        for b in self.blocks:
            a = b._window.acc
            d_conv_w, d_norm_w, d_norm_b, d_fc1_w, d_fc2_w, d_fc2_b, d_gamma = ops.unfold_block_grads(
                a["dw1f"], a["db1f"], a["dw2f"], a["db2f"], a["dwt"], b.norm.weight.detach(),
                b.norm.bias.detach(), b.mlp.fc1.weight.detach(), b.mlp.fc2.weight.detach(),
                b.mlp.fc2.bias.detach(), b.gamma.detach())
            for prm, g in ((b.conv_dw.weight, d_conv_w), (b.conv_dw.bias, a["dbias"].clone()),
                           (b.norm.weight, d_norm_w), (b.norm.bias, d_norm_b),
                           (b.mlp.fc1.weight, d_fc1_w), (b.mlp.fc1.bias, a["db1f"].clone()),
                           (b.mlp.fc2.weight, d_fc2_w), (b.mlp.fc2.bias, d_fc2_b),
                           (b.gamma, d_gamma)):
                if prm.requires_grad:
                    prm.grad = g if prm.grad is None else prm.grad + g
        self.end()

    def detach(self):
        for b in self.blocks:
            if b._window is not None and b._window.owner is self:
                b._window = None
        self.active = False


def _numel(shape):
    n = 1
    for d in shape:
        n *= d
    return n


def swap_blocks(model: nn.Module) -> int:
    """Replace every timm-style ConvNeXt block in `model.stages[i].blocks[j]` by a
    ConvNeXtBlockB200 holding the same weights.  Returns the number of blocks swapped."""
    n = 0
    for stage in model.stages:
        for j, blk in enumerate(stage.blocks):
            if isinstance(blk, ConvNeXtBlockB200):
                continue
            if not all(hasattr(blk, a) for a in ("conv_dw", "norm", "mlp", "gamma")):
                raise _lib.CnxError(f"block {type(blk).__name__} does not expose timm's "
                                    "ConvNeXtBlock attributes")
            dim = blk.conv_dw.out_channels
            new = ConvNeXtBlockB200(dim).to(blk.conv_dw.weight.device).load_from(blk)
            new.train(blk.training)
            stage.blocks[j] = new
            n += 1
    return n
