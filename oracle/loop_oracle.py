"""Oracle: the reference training loop on CPU in plain PyTorch.  TEST INFRASTRUCTURE and the
`cpu_baseline` / `--impl reference` arm of bench.py (the one place outside tests that may run it).

Restates /root/reference/train_advanced_v4.py: `EMA` (:62-94), the non-AMP branch of
`train_one_epoch` (:240-258; on CPU the reference passes scaler=None so this is the branch that
runs), `create_model` (:484-502, with the oracle block standing in for timm's) and the per-sample
MixUp/CutMix of dataset.py through oracle/mix_oracle.py.
"""
from __future__ import annotations

import numpy as np
import torch

from . import block_oracle, mix_oracle


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


def build_oracle_model(name: str = "convnext_base", num_classes: int = 44, ls_init_value=1e-6,
                       deep_supervision: bool = False):
    # skeleton / wrapper only: on CPU tensors every module of these classes is the plain PyTorch op
    from image_classification_b200.convnext import create_convnext, DeepSupervisionModelB200
    base = create_convnext(name, num_classes, block_cls=block_oracle.OracleConvNeXtBlock,
                           ls_init_value=ls_init_value)
    # config.py:22 use_deep_supervision = True -> DeepSupervisionModel(base, num_classes) (:494-495)
    return DeepSupervisionModelB200(base, num_classes) if deep_supervision else base


def reference_cpu_step(model, optimizer, scheduler, ema, images, labels, *, accum: int,
                       clip: float, label_smoothing: float, num_classes: int, rng) -> float:
    """One optimizer step = `accum` micro-batches; images [accum, b, 3, R, R] fp32 on CPU."""
    criterion = torch.nn.CrossEntropyLoss(label_smoothing=label_smoothing)   # :626
    last = 0.0
    for i in range(accum):
        mixed, targets, _ = mix_oracle.mix_dataset_ref(
            images[i].numpy(), labels[i].numpy(), mixup_alpha=0.2, cutmix_alpha=1.0, prob=0.5,
            num_classes=num_classes, rng=rng)                                 # :593-599
        out = model(torch.from_numpy(mixed))                                  # :240
        if isinstance(out, list):
            out = out[0]                                                      # :241-242
        loss = criterion(out, torch.from_numpy(targets)) / accum              # :243-244
        loss.backward()                                                       # :245
        last = float(loss.item())                                             # :268
    if clip > 0:
        torch.nn.utils.clip_grad_norm_(model.parameters(), clip)              # :249
    optimizer.step()                                                          # :250
    optimizer.zero_grad()                                                     # :251
    if scheduler is not None:
        scheduler.step()                                                      # :254
    if ema is not None:
        ema.update()                                                          # :257
    return last
