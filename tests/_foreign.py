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
