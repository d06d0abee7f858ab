"""Oracle: pure-PyTorch restatement of timm's ConvNeXtBlock and of the loss / EMA glue.

TEST INFRASTRUCTURE (see oracle/__init__.py).  parity unpinned for the block: the arithmetic
lives in the un-vendored `timm` package; call sites in the reference are
/root/reference/train_advanced_v4.py:488-492 (create_model) and :134-135 (stage(x)).  The same
math is torchvision's `CNBlock` (torchvision/models/convnext.py:39-67), used as cross-check.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn

LN_EPS = 1e-6


def block_forward_ref(x, conv_w, conv_b, norm_w, norm_b, fc1_w, fc1_b, fc2_w, fc2_b, gamma,
                      eps: float = LN_EPS):
    """timm ConvNeXtBlock.forward (conv_mlp=False), NCHW in / NCHW out, any float dtype.

    shortcut = x; x = conv_dw(x); x = x.permute(0,2,3,1); x = norm(x); x = fc1(x); x = gelu(x);
    x = fc2(x); x = x.permute(0,3,1,2); x = x * gamma[None,:,None,None]; return x + shortcut
    """
    C = x.shape[1]
    y = F.conv2d(x, conv_w, conv_b, padding=3, groups=C)
    y = y.permute(0, 2, 3, 1)
    y = F.layer_norm(y, (C,), norm_w, norm_b, eps)
    y = F.linear(y, fc1_w, fc1_b)
    y = F.gelu(y)                      # exact erf form (approximate='none')
    y = F.linear(y, fc2_w, fc2_b)
    y = y.permute(0, 3, 1, 2)
    y = y * gamma.reshape(1, -1, 1, 1)
    return y + x


def dwconv_ln_ref(x_nhwc, conv_w, conv_b, eps: float = LN_EPS, amp_round=None):
    """Reference for cnx_dwconv_ln_fwd: x [B,H,W,C] -> (xhat, mean, rstd) in x's dtype,
    LayerNorm WITHOUT affine (the affine is folded into fc1 in the product).

    amp_round=torch.bfloat16 restates the reference's autocast numerics: under
    torch.amp.autocast (/root/reference/train_advanced_v4.py:217) conv2d returns the half type and
    layer_norm then runs in fp32 ON THAT ROUNDED tensor."""
    B, H, W, C = x_nhwc.shape
    y = F.conv2d(x_nhwc.permute(0, 3, 1, 2), conv_w, conv_b, padding=3, groups=C)
    y = y.permute(0, 2, 3, 1)
    if amp_round is not None:
        y = y.to(amp_round).to(x_nhwc.dtype)
    mean = y.mean(-1)
    var = y.var(-1, unbiased=False)
    rstd = torch.rsqrt(var + eps)
    xhat = (y - mean[..., None]) * rstd[..., None]
    return xhat, mean.reshape(-1), rstd.reshape(-1)


class _Mlp(nn.Module):
    def __init__(self, dim, hidden):
        super().__init__()
        self.fc1 = nn.Linear(dim, hidden)
        self.act = nn.GELU()
        self.fc2 = nn.Linear(hidden, dim)


class OracleConvNeXtBlock(nn.Module):
    """timm-named ConvNeXt block in plain PyTorch (CPU or GPU, any dtype)."""

    def __init__(self, dim: int, ls_init_value: float | None = 1e-6):
        super().__init__()
        self.conv_dw = nn.Conv2d(dim, dim, 7, padding=3, groups=dim)
        self.norm = nn.LayerNorm(dim, eps=LN_EPS)
        self.mlp = _Mlp(dim, 4 * dim)
        init = 1.0 if ls_init_value is None else ls_init_value
        self.gamma = nn.Parameter(init * torch.ones(dim))

    def forward(self, x):
        return block_forward_ref(x, self.conv_dw.weight, self.conv_dw.bias, self.norm.weight,
                                 self.norm.bias, self.mlp.fc1.weight, self.mlp.fc1.bias,
                                 self.mlp.fc2.weight, self.mlp.fc2.bias, self.gamma)


def soft_ce_ref(logits, targets, label_smoothing: float):
    """loss = mean_b( -sum_c ((1-eps) t_bc + eps/C) log_softmax(z_b)_c )   (SURVEY.md §8 a4);
    what nn.CrossEntropyLoss(label_smoothing=eps) computes for probability targets
    (/root/reference/train_advanced_v4.py:626, :221)."""
    C = logits.shape[1]
    t = targets * (1.0 - label_smoothing) + label_smoothing / C
    return -(t * F.log_softmax(logits, dim=1)).sum(1).mean()


def ema_update_ref(shadow: torch.Tensor, param: torch.Tensor, decay: float) -> torch.Tensor:
    """EMA.update for one tensor (/root/reference/train_advanced_v4.py:79)."""
    return (1.0 - decay) * param + decay * shadow


def convert_torchvision_convnext(tv_state: dict) -> dict:
    """torchvision ConvNeXt state_dict -> timm-style names (SURVEY.md §8c key map)."""
    out = {}
    stage_of = {1: 0, 3: 1, 5: 2, 7: 3}
    down_of = {2: 1, 4: 2, 6: 3}
    sub = {"0": "conv_dw", "2": "norm", "3": "mlp.fc1", "5": "mlp.fc2"}
    for k, v in tv_state.items():
        p = k.split(".")
        if p[0] == "features":
            i = int(p[1])
            if i == 0:
                out[f"stem.{p[2]}.{p[3]}"] = v
            elif i in stage_of:
                s, j = stage_of[i], p[2]
                if p[3] == "layer_scale":
                    out[f"stages.{s}.blocks.{j}.gamma"] = v.reshape(-1)
                else:  # block.<idx>.<weight|bias>
                    out[f"stages.{s}.blocks.{j}.{sub[p[4]]}.{p[5]}"] = v
            elif i in down_of:
                out[f"stages.{down_of[i]}.downsample.{p[2]}.{p[3]}"] = v
        elif p[0] == "classifier":
            name = {"0": "head.norm", "2": "head.fc"}[p[1]]
            out[f"{name}.{p[2]}"] = v
    return out
