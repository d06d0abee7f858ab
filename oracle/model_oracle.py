"""Oracle: the model `create_model` builds, in plain PyTorch.  TEST INFRASTRUCTURE ONLY.

Restates, with nothing imported from the product package:
  * the object `timm.create_model('convnext_base', pretrained=False, num_classes=44)` returns
    (/root/reference/train_advanced_v4.py:488-492; timm is un-vendored, requirements.txt:3), with
    the attribute layout the reference touches (`stem`, `stages[i].downsample|blocks`,
    `head.{global_pool,norm,flatten,drop,fc}`, :106-111, :131-143) — pinned to torchvision's
    architecture-identical `convnext_tiny` through the SURVEY.md §8c key map
    (tests/test_host_logic.py::test_skeleton_equals_torchvision_convnext);
  * `DeepSupervisionModel` (/root/reference/train_advanced_v4.py:96-152): three auxiliary
    pool + flatten + Linear classifiers on the outputs of stages 2-4, forward child by child,
    returns `[main_out] + aux_outs`.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn

from .block_oracle import OracleConvNeXtBlock

CFGS = {
    "convnext_tiny": dict(depths=(3, 3, 9, 3), dims=(96, 192, 384, 768)),
    "convnext_small": dict(depths=(3, 3, 27, 3), dims=(96, 192, 384, 768)),
    "convnext_base": dict(depths=(3, 3, 27, 3), dims=(128, 256, 512, 1024)),
}


class OracleLayerNorm2d(nn.LayerNorm):
    """LayerNorm over the channel dimension of an NCHW map (timm's LayerNorm2d)."""

    def forward(self, x):
        y = F.layer_norm(x.permute(0, 2, 3, 1), self.normalized_shape, self.weight, self.bias, self.eps)
        return y.permute(0, 3, 1, 2)


class _Stage(nn.Module):
    def __init__(self, cin, cout, depth, stride, ls_init_value):
        super().__init__()
        if cin != cout or stride > 1:
            self.downsample = nn.Sequential(OracleLayerNorm2d(cin, eps=1e-6),
                                            nn.Conv2d(cin, cout, kernel_size=stride, stride=stride))
        else:
            self.downsample = nn.Identity()
        self.blocks = nn.Sequential(*[OracleConvNeXtBlock(cout, ls_init_value) for _ in range(depth)])

    def forward(self, x):
        return self.blocks(self.downsample(x))


class _Head(nn.Module):
    def __init__(self, dim, num_classes):
        super().__init__()
        self.global_pool = nn.AdaptiveAvgPool2d(1)
        self.norm = OracleLayerNorm2d(dim, eps=1e-6)
        self.flatten = nn.Flatten(1)
        self.drop = nn.Dropout(0.0)
        self.fc = nn.Linear(dim, num_classes)

    def forward(self, x):
        return self.fc(self.drop(self.flatten(self.norm(self.global_pool(x)))))


class OracleConvNeXt(nn.Module):
    def __init__(self, depths, dims, num_classes=44, in_chans=3, ls_init_value=1e-6):
        super().__init__()
        self.stem = nn.Sequential(nn.Conv2d(in_chans, dims[0], kernel_size=4, stride=4),
                                  OracleLayerNorm2d(dims[0], eps=1e-6))
        stages, prev = [], dims[0]
        for i, (d, c) in enumerate(zip(depths, dims)):
            stages.append(_Stage(prev, c, d, 1 if i == 0 else 2, ls_init_value))
            prev = c
        self.stages = nn.Sequential(*stages)
        self.head = _Head(prev, num_classes)
        for m in self.modules():            # trunc-normal(0.02) weights, zero bias (timm / torchvision)
            if isinstance(m, (nn.Conv2d, nn.Linear)):
                nn.init.trunc_normal_(m.weight, std=0.02)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)

    def forward(self, x):
        return self.head(self.stages(self.stem(x)))


class OracleDeepSupervision(nn.Module):
    """/root/reference/train_advanced_v4.py:96-152."""

    def __init__(self, base_model, num_classes):
        super().__init__()
        self.base_model = base_model
        channels = [st.blocks[-1].conv_dw.out_channels for st in base_model.stages[1:]]   # :105-111
        self.aux_classifiers = nn.ModuleList([
            nn.Sequential(nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(c, num_classes))
            for c in channels])                                                              # :118-124

    def forward(self, x):
        features = []
        x = self.base_model.stem(x)                                                          # :131
        for i, stage in enumerate(self.base_model.stages):                                   # :134-137
            x = stage(x)
            if i > 0:
                features.append(x)
        h = self.base_model.head
        main_out = h.fc(h.flatten(h.norm(h.global_pool(x))))                                 # :140-143
        return [main_out] + [clf(f) for f, clf in zip(features, self.aux_classifiers)]      # :146-151


def build_oracle_model(name="convnext_base", num_classes=44, ls_init_value=1e-6,
                       deep_supervision=False, depths=None, dims=None):
    cfg = CFGS.get(name, {})
    base = OracleConvNeXt(depths or cfg["depths"], dims or cfg["dims"], num_classes,
                          ls_init_value=ls_init_value)
    return OracleDeepSupervision(base, num_classes) if deep_supervision else base
