"""A timm-style ConvNeXt assembled from plain torch modules (what `timm.create_model` would hand
over), for the swap_blocks / swap_trunk tests."""
import torch


class _TimmLikeLN2d(torch.nn.LayerNorm):
    """timm's LayerNorm2d as a third party would have built it (no libcnx_b200 path)."""

    def forward(self, x):
        y = torch.nn.functional.layer_norm(x.permute(0, 2, 3, 1), self.normalized_shape, self.weight,
                                           self.bias, self.eps)
        return y.permute(0, 3, 1, 2)


def _foreign_convnext(depths, dims):
    """ConvNeXtTimmLike whose stem / downsample / head modules are plain torch ones."""
    from image_classification_b200.convnext import ConvNeXtTimmLike
    from oracle import block_oracle
    m = ConvNeXtTimmLike(depths, dims, 44, block_cls=block_oracle.OracleConvNeXtBlock)

    def conv_of(c):
        n = torch.nn.Conv2d(c.in_channels, c.out_channels, c.kernel_size, c.stride)
        n.load_state_dict(c.state_dict())
        return n

    def ln_of(l):
        n = _TimmLikeLN2d(l.normalized_shape[0], eps=l.eps)
        n.load_state_dict(l.state_dict())
        return n

    m.stem = torch.nn.Sequential(conv_of(m.stem[0]), ln_of(m.stem[1]))
    for st in m.stages:
        if not isinstance(st.downsample, torch.nn.Identity):
            st.downsample = torch.nn.Sequential(ln_of(st.downsample[0]), conv_of(st.downsample[1]))
    m.head.global_pool = torch.nn.AdaptiveAvgPool2d(1)
    m.head.norm = ln_of(m.head.norm)
    return m


# ----------------------------------------------------------------------------------------------
# timm-style EfficientNet pieces (timm is not installable here).  Attribute names, module nesting
# and state_dict keys follow timm's `efficientnet_b2` [upstream]: BatchNorm and activation fused in
# one `BatchNormAct2d` (an nn.BatchNorm2d subclass with `.drop` / `.act`), squeeze-excite with
# `conv_reduce` / `act1` / `conv_expand` / `gate`, blocks `DepthwiseSeparableConv`
# (conv_dw, bn1, se, conv_pw, bn2) and `InvertedResidual` (conv_pw, bn1, conv_dw, bn2, se,
# conv_pwl, bn3) that are plain modules, not nn.Sequential.
class TimmLikeBatchNormAct2d(torch.nn.BatchNorm2d):
    def __init__(self, num_features, apply_act=True):
        super().__init__(num_features, eps=1e-5, momentum=0.1)
        self.drop = torch.nn.Identity()
        self.act = torch.nn.SiLU(inplace=True) if apply_act else torch.nn.Identity()

    def forward(self, x):
        return self.act(self.drop(super().forward(x)))


class TimmLikeSqueezeExcite(torch.nn.Module):
    def __init__(self, chs, rd):
        super().__init__()
        self.conv_reduce = torch.nn.Conv2d(chs, rd, 1, bias=True)
        self.act1 = torch.nn.SiLU(inplace=True)
        self.conv_expand = torch.nn.Conv2d(rd, chs, 1, bias=True)
        self.gate = torch.nn.Sigmoid()

    def forward(self, x):
        s = x.mean((2, 3), keepdim=True)
        return x * self.gate(self.conv_expand(self.act1(self.conv_reduce(s))))


class TimmLikeDepthwiseSeparableConv(torch.nn.Module):
    def __init__(self, cin, cout, k=3, stride=1):
        super().__init__()
        self.has_skip = stride == 1 and cin == cout
        self.conv_dw = torch.nn.Conv2d(cin, cin, k, stride, k // 2, groups=cin, bias=False)
        self.bn1 = TimmLikeBatchNormAct2d(cin)
        self.se = TimmLikeSqueezeExcite(cin, max(1, cin // 4))
        self.conv_pw = torch.nn.Conv2d(cin, cout, 1, bias=False)
        self.bn2 = TimmLikeBatchNormAct2d(cout, apply_act=False)

    def forward(self, x):
        y = self.bn2(self.conv_pw(self.se(self.bn1(self.conv_dw(x)))))
        return y + x if self.has_skip else y


class TimmLikeInvertedResidual(torch.nn.Module):
    def __init__(self, cin, cout, k=3, stride=1, exp=6):
        super().__init__()
        mid = cin * exp
        self.has_skip = stride == 1 and cin == cout
        self.conv_pw = torch.nn.Conv2d(cin, mid, 1, bias=False)
        self.bn1 = TimmLikeBatchNormAct2d(mid)
        self.conv_dw = torch.nn.Conv2d(mid, mid, k, stride, k // 2, groups=mid, bias=False)
        self.bn2 = TimmLikeBatchNormAct2d(mid)
        self.se = TimmLikeSqueezeExcite(mid, max(1, cin // 4))
        self.conv_pwl = torch.nn.Conv2d(mid, cout, 1, bias=False)
        self.bn3 = TimmLikeBatchNormAct2d(cout, apply_act=False)

    def forward(self, x):
        y = self.bn1(self.conv_pw(x))
        y = self.se(self.bn2(self.conv_dw(y)))
        y = self.bn3(self.conv_pwl(y))
        return y + x if self.has_skip else y


class TimmLikeEfficientNet(torch.nn.Module):
    """conv_stem / bn1 / blocks / conv_head / bn2 / global_pool / classifier, the first stages of
    efficientnet_b2's widths (32 -> 16 -> 24 -> 48) so that a CPU test stays small."""

    def __init__(self, num_classes=44):
        super().__init__()
        self.conv_stem = torch.nn.Conv2d(3, 32, 3, 2, 1, bias=False)
        self.bn1 = TimmLikeBatchNormAct2d(32)
        self.blocks = torch.nn.Sequential(
            torch.nn.Sequential(TimmLikeDepthwiseSeparableConv(32, 16), TimmLikeDepthwiseSeparableConv(16, 16)),
            torch.nn.Sequential(TimmLikeInvertedResidual(16, 24, 3, 2), TimmLikeInvertedResidual(24, 24, 3, 1)),
            torch.nn.Sequential(TimmLikeInvertedResidual(24, 48, 5, 2), TimmLikeInvertedResidual(48, 48, 5, 1)))
        self.conv_head = torch.nn.Conv2d(48, 128, 1, bias=False)
        self.bn2 = TimmLikeBatchNormAct2d(128)
        self.global_pool = torch.nn.AdaptiveAvgPool2d(1)
        self.classifier = torch.nn.Linear(128, num_classes)

    def forward(self, x):
        x = self.bn1(self.conv_stem(x))
        x = self.bn2(self.conv_head(self.blocks(x)))
        return self.classifier(self.global_pool(x).flatten(1))
