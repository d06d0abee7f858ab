"""Oracle: the reference training loop on CPU in plain PyTorch.  TEST INFRASTRUCTURE and the
`cpu_baseline` / `--impl reference` arm of bench.py (the one place outside tests that may run it).

Restates /root/reference/train_advanced_v4.py: `EMA` (:62-94), BOTH branches of `train_one_epoch`
(:215-237 AMP, :238-258 plain), `create_model` (:484-502 through oracle/model_oracle.py) and the
per-sample MixUp/CutMix of dataset.py through oracle/mix_oracle.py.  Nothing here imports the
product package.

PINNED: `oracle/gen_golden.py::gen_loop` runs the reference's OWN `train_one_epoch`, `EMA` and
`get_cosine_schedule_with_warmup` (imported verbatim by oracle/ref_loader.py) on a tiny
list-returning model and commits parameters after every micro-batch, EMA shadows, the LR trace and
the returned mean loss as tests/golden/loop_reference.npz; tests/test_oracle_golden.py checks
`reference_loop` below against it (1e-6), tests/test_gpu_loop.py checks the product's `TrainStep`.
"""
from __future__ import annotations

import math

import numpy as np
import torch
from torch import nn

from . import mix_oracle
from .model_oracle import build_oracle_model  # noqa: F401  (re-exported for bench.py / tests)


class EMARef:
    """train_advanced_v4.py:62-94, verbatim semantics (Python loop, clone per tensor)."""

    def __init__(self, model, decay):
        self.model, self.decay, self.shadow, self.backup = model, decay, {}, {}
        for name, param in self.model.named_parameters():      # register(), :70-73
            if param.requires_grad:
                self.shadow[name] = param.data.clone()

    def update(self):                                           # :75-80
        for name, param in self.model.named_parameters():
            if param.requires_grad:
                new_average = (1.0 - self.decay) * param.data + self.decay * self.shadow[name]
                self.shadow[name] = new_average.clone()


def cosine_warmup_lambda(num_warmup_steps, num_training_steps, min_lr=1e-6):
    """:183-191."""
    def lr_lambda(step):
        if step < num_warmup_steps:
            return float(step) / float(max(1, num_warmup_steps))
        progress = float(step - num_warmup_steps) / float(max(1, num_training_steps - num_warmup_steps))
        return max(min_lr, 0.5 * (1.0 + math.cos(math.pi * progress)))
    return lr_lambda


class TinyListModel(nn.Module):
    """The fixture model of gen_loop: flatten -> Linear -> GELU -> LayerNorm -> Linear, returning
    `[main, aux]` like DeepSupervisionModel (:151) so the `outputs[0]` selection (:219-220) runs;
    the auxiliary head never receives a gradient, like the reference's three (:146-149)."""

    def __init__(self, in_features=3 * 8 * 8, hidden=32, num_classes=10):
        super().__init__()
        self.fc1 = nn.Linear(in_features, hidden)
        self.norm = nn.LayerNorm(hidden)
        self.fc2 = nn.Linear(hidden, num_classes)
        self.aux = nn.Linear(hidden, num_classes)

    def forward(self, x):
        h = self.norm(torch.nn.functional.gelu(self.fc1(x.flatten(1))))
        return [self.fc2(h), self.aux(h)]


def reference_loop(model, batches, criterion, optimizer, scheduler, ema, *, accum: int, clip: float,
                   amp_branch: bool, on_micro=None) -> float:
    """`train_one_epoch` over `batches` = [(images, soft labels)], restated.  amp_branch=True is the
    `scaler is not None` branch with a scaler whose scale is 1 (:215-237: the loss is NOT divided by
    `accum`); False the plain branch (:238-258).  Returns `losses.avg` (:268, weighted by batch
    size).  on_micro(i) is called after every micro-batch."""
    model.train()
    s = n = 0.0
    for batch_idx, (images, labels) in enumerate(batches):
        outputs = model(images)                                               # :218 / :240
        if isinstance(outputs, list):
            outputs = outputs[0]                                              # :219-220 / :241-242
        loss = criterion(outputs, labels)                                     # :221 / :243
        if not amp_branch:
            loss = loss / accum                                               # :244 (absent at :223)
        loss.backward()                                                       # :223 / :245
        if (batch_idx + 1) % accum == 0:                                      # :225 / :247
            if clip > 0:
                torch.nn.utils.clip_grad_norm_(model.parameters(), clip)      # :228 / :249
            optimizer.step()                                                  # :229 / :250
            optimizer.zero_grad()                                             # :231 / :251
            if scheduler is not None:
                scheduler.step()                                              # :233-234 / :253-254
            if ema is not None:
                ema.update()                                                  # :236-237 / :256-257
        s += float(loss.item()) * images.size(0)                              # :268
        n += images.size(0)
        if on_micro is not None:
            on_micro(batch_idx)
    return s / max(1.0, n)


def reference_cpu_step(model, optimizer, scheduler, ema, images, labels, *, accum: int,
                       clip: float, label_smoothing: float, num_classes: int, rng) -> float:
    """One optimizer step = `accum` micro-batches; images [accum, b, 3, R, R] fp32 on CPU.  On CPU
    the reference passes scaler=None (:432 is CUDA-only), so the plain branch runs."""
    criterion = torch.nn.CrossEntropyLoss(label_smoothing=label_smoothing)   # :626
    batches = []
    for i in range(accum):
        mixed, targets, _ = mix_oracle.mix_dataset_ref(
            images[i].numpy(), labels[i].numpy(), mixup_alpha=0.2, cutmix_alpha=1.0, prob=0.5,
            num_classes=num_classes, rng=rng)                                 # :593-599
        batches.append((torch.from_numpy(mixed), torch.from_numpy(targets)))
    return reference_loop(model, batches, criterion, optimizer, scheduler, ema, accum=accum,
                          clip=clip, amp_branch=False)
