"""Label-smoothed soft-target cross-entropy as ONE kernel (loss + dlogits).

Stands in for `nn.CrossEntropyLoss(label_smoothing=cfg.label_smoothing)` as the reference builds
and applies it (/root/reference/train_advanced_v4.py:626, :221) on the `[B, 44]` probability
targets produced by MixUp/CutMix (/root/reference/dataset.py:107,153,172).
"""
from __future__ import annotations

import torch
from torch import nn

from . import _lib


class _SoftCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits: torch.Tensor, targets: torch.Tensor, label_smoothing: float):
        if not logits.is_cuda:
            raise _lib.CnxError("SoftTargetCrossEntropyB200: logits must be CUDA (no CPU path)")
        in_dtype = logits.dtype
        z = logits.detach().float().contiguous()   # CE runs in fp32 under autocast
        t = targets.detach().to(device=z.device, dtype=torch.float32).contiguous()
        B, NC = z.shape
        if t.shape != z.shape:
            raise _lib.CnxError(f"targets {tuple(t.shape)} do not match logits {tuple(z.shape)}")
        loss = torch.empty((), device=z.device, dtype=torch.float32)
        dz = torch.empty_like(z)
        _lib.call("cnx_softce_fwd_bwd", z.data_ptr(), t.data_ptr(), loss.data_ptr(), dz.data_ptr(),
                  B, NC, float(label_smoothing), torch.cuda.current_stream(z.device).cuda_stream)
        ctx.save_for_backward(dz)
        ctx.in_dtype = in_dtype
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        (dz,) = ctx.saved_tensors
        return (dz * grad_out).to(ctx.in_dtype), None, None


class SoftTargetCrossEntropyB200(nn.Module):
    """`criterion(outputs, labels)` with the reference's semantics: probability targets `[B, C]`
    (or class indices `[B]`, turned into one-hot rows), mean reduction, label smoothing eps."""

    def __init__(self, label_smoothing: float = 0.0):
        super().__init__()
        self.label_smoothing = float(label_smoothing)

    def forward(self, logits: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
        if targets.dim() == 1:
            targets = torch.nn.functional.one_hot(targets.long(), logits.shape[1]).float()
        return _SoftCE.apply(logits, targets, self.label_smoothing)
