"""The training micro-step, in the reference's order of operations.

Mirror of `train_one_epoch` (/root/reference/train_advanced_v4.py:193-281) and of
`get_cosine_schedule_with_warmup` (:183-191) for the B200 path.  Semantics kept on purpose
(SURVEY.md H7): in the AMP branch the loss is NOT divided by the accumulation steps (:223 vs
:244); the optimizer/clip/scheduler/EMA sequence runs every `gradient_accumulation_steps`-th
batch (:225-237); the scheduler steps once per optimizer step; `min_lr` is a multiplier (:189);
`GradScaler` calls stay in place although bf16 autocast makes them near no-ops (BASELINE mandates
bf16 where the reference's autocast default is fp16).  Differences, also on purpose: errors are
raised, not swallowed per batch (:277-279 would hide a failing kernel), and `loss.item()` (:268,
a device sync per micro-step) is replaced by a device-side running sum read once per epoch.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch


def cosine_schedule_with_warmup_lambda(num_warmup_steps: int, num_training_steps: int,
                                       min_lr: float = 1e-6):
    """The `lr_lambda` of the reference schedule (train_advanced_v4.py:185-189)."""
    def lr_lambda(current_step):
        if current_step < num_warmup_steps:
            return float(current_step) / float(max(1, num_warmup_steps))
        progress = float(current_step - num_warmup_steps) / float(
            max(1, num_training_steps - num_warmup_steps))
        return max(min_lr, 0.5 * (1.0 + math.cos(math.pi * progress)))
    return lr_lambda


def get_cosine_schedule_with_warmup(optimizer, num_warmup_steps, num_training_steps, min_lr=1e-6):
    return torch.optim.lr_scheduler.LambdaLR(
        optimizer, cosine_schedule_with_warmup_lambda(num_warmup_steps, num_training_steps, min_lr))


@dataclass
class StepConfig:
    """The fields of the reference `Config` the step loop reads (config.py:31-48)."""

    gradient_accumulation_steps: int = 2
    gradient_clip_val: float = 1.0
    num_classes: int = 44
    amp_dtype: torch.dtype = torch.bfloat16


class TrainStep:
    """One object per (model, optimizer): call `micro_step(images, labels, batch_idx)` for
    every batch, exactly like the body of the reference's `for batch_idx, ...` loop."""

    def __init__(self, model, criterion, optimizer, scheduler=None, scaler=None, ema=None,
                 cfg: StepConfig | None = None, grad_reducer=None, use_grad_window: bool = True,
                 use_fused_optim: bool | None = None):
        self.model, self.criterion, self.optimizer = model, criterion, optimizer
        self.scheduler, self.scaler, self.ema = scheduler, scaler, ema
        self.cfg = cfg or StepConfig()
        self.grad_reducer = grad_reducer          # ddp.BucketedGradReducer or None
        self.params = [p for p in model.parameters() if p.requires_grad]
        # with accumulation > 1 the B200 blocks keep folded-parameter gradients in one fp32 window
        # and run the fold adjoint once per optimizer step (block.GradWindow); GradScaler's scale is
        # constant inside a window (it only changes in scaler.update(), :230)
        # lines :226-230 and :236-237 as two launches (optim_step.FusedOptimStep) whenever the
        # optimizer is the reference's AdamW over CUDA fp32 parameters; otherwise torch's own calls
        self.fused = None
        if use_fused_optim is None:
            use_fused_optim = (isinstance(optimizer, torch.optim.AdamW) and len(self.params) > 0 and
                               all(p.is_cuda and p.dtype == torch.float32 for p in self.params) and
                               not any(g.get("amsgrad") or g.get("maximize")
                                       for g in optimizer.param_groups))
        if use_fused_optim:
            from .optim_step import FusedOptimStep
            self.fused = FusedOptimStep(model, optimizer, scaler, ema, self.cfg.gradient_clip_val)
        from .block import ConvNeXtBlockB200
        self.blocks = [m for m in model.modules() if isinstance(m, ConvNeXtBlockB200)]
        self.window = None
        if self.cfg.gradient_accumulation_steps > 1 and use_grad_window:
            from .block import GradWindow
            w = GradWindow(model)
            self.window = w if w.blocks else None

    def micro_step(self, images: torch.Tensor, labels: torch.Tensor, batch_idx: int) -> torch.Tensor:
        boundary = (batch_idx + 1) % self.cfg.gradient_accumulation_steps == 0     # :225
        loss = self.forward_backward(images, labels, batch_idx)
        if boundary:
            self.optimizer_section()
        return loss

    def forward_backward(self, images: torch.Tensor, labels: torch.Tensor, batch_idx: int) -> torch.Tensor:
        """:215-223 / :238-245 — forward, criterion, (scaled) backward of one micro-batch."""
        cfg = self.cfg
        boundary = (batch_idx + 1) % cfg.gradient_accumulation_steps == 0     # :225
        if self.grad_reducer is not None:
            self.grad_reducer.set_sync(boundary)       # all-reduce only on the boundary micro-step
        if self.window is not None:
            if batch_idx % cfg.gradient_accumulation_steps == 0:
                self.window.begin()
            self.window.set_boundary(boundary)
        if self.blocks and images.is_cuda and batch_idx % cfg.gradient_accumulation_steps == 0:
            # the parameters only change in the optimizer section: fold every block's weights for
            # this window now, on the side stream, under the stem
            from .block import prefold_blocks
            prefold_blocks(self.blocks, images.device)
        if self.scaler is not None:
            with torch.autocast("cuda", dtype=cfg.amp_dtype):                 # :217
                outputs = self.model(images)                                   # :218
                if isinstance(outputs, (list, tuple)):
                    outputs = outputs[0]                                       # :219-220
                loss = self.criterion(outputs, labels)                         # :221
            self.scaler.scale(loss).backward()                                 # :223 (no /accum)
        else:
            outputs = self.model(images)                                       # :240
            if isinstance(outputs, (list, tuple)):
                outputs = outputs[0]                                           # :241-242
            loss = self.criterion(outputs, labels)                             # :243
            loss = loss / cfg.gradient_accumulation_steps                      # :244
            loss.backward()                                                    # :245
        return loss.detach()

    def optimizer_section(self, reduce_grads: bool = True) -> None:
        """:226-237 / :248-257 — what runs on the boundary micro-batch after backward.
        reduce_grads=False: the data-parallel exchange already ran (inside a replayed graph)."""
        cfg = self.cfg
        if reduce_grads and self.grad_reducer is not None:
            self.grad_reducer.finish()
        if self.fused is not None:
            self.fused.step()                                                  # :226-230, :236-237
        elif self.scaler is not None:
            self.scaler.unscale_(self.optimizer)                               # :226
            if cfg.gradient_clip_val > 0:
                torch.nn.utils.clip_grad_norm_(self.params, cfg.gradient_clip_val)   # :228
            self.scaler.step(self.optimizer)                                   # :229
            self.scaler.update()                                               # :230
        else:
            if cfg.gradient_clip_val > 0:
                torch.nn.utils.clip_grad_norm_(self.params, cfg.gradient_clip_val)   # :249
            self.optimizer.step()                                              # :250
        self.optimizer.zero_grad()                                             # :231 / :251
        if self.scheduler is not None:
            self.scheduler.step()                                              # :233-234
        if self.ema is not None and self.fused is None:
            self.ema.update()                                                  # :236-237
        if self.window is not None:
            self.window.end()

    def close(self) -> None:
        """End of the loop: gradients of an odd trailing micro-batch still held by the
        accumulation window go to `.grad` (where the reference leaves them, :225-231), the blocks
        are released from the window."""
        if self.window is not None:
            self.window.flush()
            self.window.detach()
            self.window = None


def train_one_epoch(model, train_loader, criterion, optimizer, scheduler, device, cfg,
                    scaler=None, ema=None, grad_reducer=None):
    """Same call signature and return value `(mean loss, top1)` as the reference function; top1
    is 0.0 for soft labels, as in the reference (:260-265 only scores hard labels)."""
    model.train()
    step_cfg = StepConfig(gradient_accumulation_steps=cfg.gradient_accumulation_steps,
                          gradient_clip_val=cfg.gradient_clip_val,
                          num_classes=getattr(cfg, "num_classes", 44))
    step = TrainStep(model, criterion, optimizer, scheduler, scaler, ema, step_cfg, grad_reducer)
    loss_sum = torch.zeros((), device=device, dtype=torch.float32)
    count = 0
    for batch_idx, (images, labels) in enumerate(train_loader):
        images = images.to(device, non_blocking=True)                          # :211
        labels = labels.to(device, non_blocking=True)                          # :213
        loss = step.micro_step(images, labels, batch_idx)
        loss_sum += loss.float() * images.size(0)                              # AverageMeter, :268
        count += images.size(0)
    step.close()
    return (float(loss_sum.item()) / max(1, count)), 0.0
