"""Batch-level MixUp/CutMix: host planner (integers bit-exact with the reference RNG order) +
one fused device kernel.

Mirrors `MixUpCutMixDataset` (/root/reference/dataset.py:70-190): same constructor arguments,
same per-sample RNG call order (`random()` gate :167, `randint(len)` partner :175, `random()`
mode :185, `beta` :83/:112, `randint(W)` then `randint(H)` :135-136), same box clipping
(:138-141), same lambda re-adjustment (:152).  The reference mixes one sample at a time on CPU
DataLoader workers and draws the partner from the whole dataset; here the "dataset" is the batch
already resident in HBM (partners are drawn inside the local micro-batch).
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib

MODE_PASS, MODE_MIXUP, MODE_CUTMIX = 0, 1, 2

PLAN_DTYPE = np.dtype([
    ("mode", np.int32), ("partner", np.int32), ("lam", np.float32), ("one_minus_lam", np.float32),
    ("bbx1", np.int32), ("bby1", np.int32), ("bbx2", np.int32), ("bby2", np.int32),
])
assert PLAN_DTYPE.itemsize == ctypes.sizeof(_lib.MixPlanEntry) == 32


def plan_mix_batch(batch_size: int, height: int, width: int, *, mixup_alpha: float = 1.0,
                   cutmix_alpha: float = 1.0, prob: float = 0.5, rng=None) -> np.ndarray:
    """Draw the per-sample mixing plan with the reference's RNG call order.

    `rng` is anything with legacy `random()/randint()/beta()` (the `np.random` module itself, which
    is what the reference uses, or a `np.random.RandomState`).  Returns a structured array of
    `PLAN_DTYPE`; `lam` is the float64 value rounded to float32 exactly as `tensor * lam` does.
    """
    rng = np.random if rng is None else rng
    plan = np.zeros(batch_size, dtype=PLAN_DTYPE)
    H, W = height, width
    for i in range(batch_size):
        e = plan[i]
        if rng.random() > prob:                      # dataset.py:167
            e["mode"], e["partner"], e["lam"], e["one_minus_lam"] = MODE_PASS, i, 1.0, 0.0
            continue
        e["partner"] = rng.randint(batch_size)       # dataset.py:175
        if rng.random() < 0.5 and mixup_alpha > 0:   # dataset.py:185
            lam = rng.beta(mixup_alpha, mixup_alpha)  # dataset.py:83
            e["mode"] = MODE_MIXUP
        else:
            lam = rng.beta(cutmix_alpha, cutmix_alpha) if cutmix_alpha > 0 else 1  # :112-114
            cut_rat = np.sqrt(1.0 - lam)             # dataset.py:131
            cut_w = int(W * cut_rat)
            cut_h = int(H * cut_rat)
            cx = rng.randint(W)                      # dataset.py:135 (x first, then y)
            cy = rng.randint(H)
            bbx1 = int(np.clip(cx - cut_w // 2, 0, W))
            bby1 = int(np.clip(cy - cut_h // 2, 0, H))
            bbx2 = int(np.clip(cx + cut_w // 2, 0, W))
            bby2 = int(np.clip(cy + cut_h // 2, 0, H))
            lam = 1 - ((bbx2 - bbx1) * (bby2 - bby1) / (W * H))   # dataset.py:152
            e["mode"] = MODE_CUTMIX
            e["bbx1"], e["bby1"], e["bbx2"], e["bby2"] = bbx1, bby1, bbx2, bby2
        e["lam"] = np.float32(lam)
        e["one_minus_lam"] = np.float32(1 - lam)     # the subtraction happens in float64
    return plan


def mix_batch(images: torch.Tensor, labels: torch.Tensor, plan: np.ndarray, num_classes: int,
              on_value: float = 1.0, off_value: float = 0.0, out: torch.Tensor | None = None,
              targets_out: torch.Tensor | None = None):
    """Apply `plan` on the device: images [B,CH,H,W] fp32 (NCHW), labels [B] int64 ->
    (mixed images, soft targets [B,num_classes] fp32).  on/off: values of the (smoothed) one-hot
    rows that are blended (1/0 = the reference's plain one-hot).  out / targets_out: existing
    result buffers (the static inputs of a CUDA-graph step, graphed.GraphedTrainStep)."""
    if not images.is_cuda:
        raise _lib.CnxError("mix_batch: images must be a CUDA tensor (no CPU path)")
    if images.dtype != torch.float32 or not images.is_contiguous():
        raise _lib.CnxError("mix_batch: images must be contiguous float32 NCHW")
    B, CH, H, W = images.shape
    if plan.dtype != PLAN_DTYPE or plan.shape != (B,):
        raise _lib.CnxError("mix_batch: plan must come from plan_mix_batch for this batch size")
    labels = labels.to(device=images.device, dtype=torch.int64).contiguous()
    plan_dev = torch.from_numpy(plan.view(np.uint8).reshape(-1).copy()).to(images.device,
                                                                            non_blocking=True)
    if out is None:
        out = torch.empty_like(images)
    elif (out.shape != images.shape or out.dtype != torch.float32 or not out.is_contiguous()
          or out.device != images.device or out.data_ptr() == images.data_ptr()):
        raise _lib.CnxError("mix_batch: `out` must be a distinct contiguous float32 tensor shaped "
                            "like the images")
    if targets_out is None:
        targets = torch.empty(B, num_classes, device=images.device, dtype=torch.float32)
    else:
        targets = targets_out
        if (tuple(targets.shape) != (B, num_classes) or targets.dtype != torch.float32
                or not targets.is_contiguous() or targets.device != images.device):
            raise _lib.CnxError(f"mix_batch: `targets_out` must be contiguous float32 [{B}, {num_classes}]")
    stream = torch.cuda.current_stream(images.device).cuda_stream
    if on_value == 1.0 and off_value == 0.0:
        _lib.call("cnx_mix_batch", images.data_ptr(), labels.data_ptr(), plan_dev.data_ptr(),
                  out.data_ptr(), targets.data_ptr(), B, CH, H, W, num_classes, stream)
    else:
        _lib.call("cnx_mix_batch_smooth", images.data_ptr(), labels.data_ptr(), plan_dev.data_ptr(),
                  out.data_ptr(), targets.data_ptr(), B, CH, H, W, num_classes,
                  float(np.float32(on_value)), float(np.float32(off_value)), stream)
    return out, targets


@dataclass
class BatchMixUpCutMix:
    """Drop-in for the role of `MixUpCutMixDataset`, applied to a whole device-resident batch."""

    mixup_alpha: float = 1.0
    cutmix_alpha: float = 1.0
    prob: float = 0.5
    num_classes: int | None = None

    def __call__(self, images: torch.Tensor, labels: torch.Tensor, rng=None, out=None,
                 targets_out=None):
        B, _, H, W = images.shape
        plan = plan_mix_batch(B, H, W, mixup_alpha=self.mixup_alpha, cutmix_alpha=self.cutmix_alpha,
                              prob=self.prob, rng=rng)
        return mix_batch(images, labels, plan, self.num_classes, out=out, targets_out=targets_out)


def plan_timm_batch_mixup(batch_size: int, height: int, width: int, *, mixup_alpha: float = 1.0,
                          cutmix_alpha: float = 0.0, prob: float = 1.0, switch_prob: float = 0.5,
                          correct_lam: bool = True, rng=None) -> np.ndarray:
    """timm `Mixup(mode='batch')` parameters as a plan: ONE lambda / mode / box for the whole batch,
    partner = the flipped batch (`x.flip(0)`).  RNG call order of timm.data.mixup
    (`_params_per_batch`, `rand_bbox`): rand() gate, [rand() switch], beta, [randint(H) THEN
    randint(W)].  timm is not vendored in the reference (requirements.txt:3): restated from the
    upstream source the reference imports at previous/V2-convbase/dataset.py:10,119-129."""
    rng = np.random if rng is None else rng
    if batch_size % 2 != 0:
        raise _lib.CnxError("timm batch Mixup: batch size must be even")       # timm asserts this
    lam, use_cutmix = 1.0, False
    if (mixup_alpha > 0 or cutmix_alpha > 0) and rng.rand() < prob:
        if mixup_alpha > 0.0 and cutmix_alpha > 0.0:
            use_cutmix = rng.rand() < switch_prob
            lam_mix = rng.beta(cutmix_alpha, cutmix_alpha) if use_cutmix else \
                rng.beta(mixup_alpha, mixup_alpha)
        elif mixup_alpha > 0.0:
            lam_mix = rng.beta(mixup_alpha, mixup_alpha)
        else:
            use_cutmix = True
            lam_mix = rng.beta(cutmix_alpha, cutmix_alpha)
        lam = float(lam_mix)
    plan = np.zeros(batch_size, dtype=PLAN_DTYPE)
    plan["partner"] = np.arange(batch_size - 1, -1, -1)
    if lam == 1.0:
        plan["mode"], plan["lam"], plan["one_minus_lam"] = MODE_PASS, 1.0, 0.0
        return plan
    if use_cutmix:
        H, W = height, width
        ratio = np.sqrt(1 - lam)
        cut_h, cut_w = int(H * ratio), int(W * ratio)
        cy = rng.randint(0, H)
        cx = rng.randint(0, W)
        yl, yh = int(np.clip(cy - cut_h // 2, 0, H)), int(np.clip(cy + cut_h // 2, 0, H))
        xl, xh = int(np.clip(cx - cut_w // 2, 0, W)), int(np.clip(cx + cut_w // 2, 0, W))
        if correct_lam:
            lam = 1.0 - (yh - yl) * (xh - xl) / float(H * W)
        plan["mode"] = MODE_CUTMIX
        plan["bbx1"], plan["bby1"], plan["bbx2"], plan["bby2"] = xl, yl, xh, yh
    else:
        plan["mode"] = MODE_MIXUP
    plan["lam"] = np.float32(lam)
    plan["one_minus_lam"] = np.float32(1.0 - lam)
    return plan


@dataclass
class TimmBatchMixup:
    """Device-side timm `Mixup(..., mode='batch')`: same constructor arguments
    as the object the reference builds at previous/V2-convbase/dataset.py:119-129; targets are
    label-smoothed one-hot rows blended with the batch lambda (`mixup_target`)."""

    mixup_alpha: float = 1.0
    cutmix_alpha: float = 0.0
    cutmix_minmax: tuple | None = None
    prob: float = 1.0
    switch_prob: float = 0.5
    mode: str = "batch"
    correct_lam: bool = True
    label_smoothing: float = 0.1
    num_classes: int = 1000

    def __post_init__(self):
        if self.mode != "batch" or self.cutmix_minmax is not None:
            raise _lib.CnxError("TimmBatchMixup: only mode='batch' without cutmix_minmax is built")

    def __call__(self, images: torch.Tensor, labels: torch.Tensor, rng=None):
        B, _, H, W = images.shape
        plan = plan_timm_batch_mixup(B, H, W, mixup_alpha=self.mixup_alpha,
                                     cutmix_alpha=self.cutmix_alpha, prob=self.prob,
                                     switch_prob=self.switch_prob, correct_lam=self.correct_lam,
                                     rng=rng)
        off = self.label_smoothing / self.num_classes
        on = 1.0 - self.label_smoothing + off
        return mix_batch(images, labels, plan, self.num_classes, on, off)
