"""Oracle: numpy restatement of MixUpCutMixDataset, sample by sample.  TEST INFRASTRUCTURE.

Follows /root/reference/dataset.py line by line (citations inline).  Pinned against the
reference's own class run on seeded inputs: see oracle/gen_golden.py and
tests/test_oracle_golden.py (bit-exact plan integers, pixels and targets).
"""
from __future__ import annotations

import numpy as np


def _onehot(label: int, num_classes: int) -> np.ndarray:
    t = np.zeros(num_classes, dtype=np.float32)      # dataset.py:102-105
    t[int(label)] = 1
    return t


def mixup_ref(data1, data2, t1, t2, alpha, num_classes, rng):
    """dataset.py:81-108.  float32 arrays times a float64 scalar: torch rounds the scalar to
    float32 first, then multiplies, then adds (three roundings)."""
    lam = rng.beta(alpha, alpha) if alpha > 0 else 1                    # :82-85
    lam32, oml32 = np.float32(lam), np.float32(1 - lam)
    data = data1 * lam32 + data2 * oml32                                # :106
    target = _onehot(t1, num_classes) * lam32 + _onehot(t2, num_classes) * oml32   # :107
    return data.astype(np.float32), target.astype(np.float32), dict(mode=1, lam=lam)


def cutmix_ref(data1, data2, t1, t2, alpha, num_classes, rng):
    """dataset.py:110-155."""
    lam = rng.beta(alpha, alpha) if alpha > 0 else 1                    # :111-114
    H, W = data1.shape[1:]                                              # :134
    cut_rat = np.sqrt(1.0 - lam)                                        # :135
    cut_w = int(W * cut_rat)
    cut_h = int(H * cut_rat)
    cx = rng.randint(W)                                                 # :139
    cy = rng.randint(H)                                                 # :140
    bbx1 = np.clip(cx - cut_w // 2, 0, W)                               # :142-145
    bby1 = np.clip(cy - cut_h // 2, 0, H)
    bbx2 = np.clip(cx + cut_w // 2, 0, W)
    bby2 = np.clip(cy + cut_h // 2, 0, H)
    data = data1.copy()                                                 # :147
    data[:, bby1:bby2, bbx1:bbx2] = data2[:, bby1:bby2, bbx1:bbx2]      # :148
    lam = 1 - ((bbx2 - bbx1) * (bby2 - bby1) / (W * H))                 # :151
    lam32, oml32 = np.float32(lam), np.float32(1 - lam)
    target = _onehot(t1, num_classes) * lam32 + _onehot(t2, num_classes) * oml32   # :152
    return data, target.astype(np.float32), dict(mode=2, lam=lam, box=(int(bbx1), int(bby1),
                                                                       int(bbx2), int(bby2)))


def mix_dataset_ref(images: np.ndarray, labels: np.ndarray, *, mixup_alpha, cutmix_alpha, prob,
                    num_classes, rng):
    """Iterate `MixUpCutMixDataset.__getitem__(i)` for i = 0..N-1 (dataset.py:157-190) over an
    in-memory dataset `images [N,CH,H,W] float32`, `labels [N]`.  Returns (mixed, targets, info)."""
    N = images.shape[0]
    out = np.empty_like(images)
    targets = np.empty((N, num_classes), dtype=np.float32)
    info = []
    for i in range(N):
        data1, t1 = images[i], labels[i]                                # :158
        if rng.random() > prob:                                         # :167
            out[i], targets[i] = data1, _onehot(t1, num_classes)        # :168-172
            info.append(dict(mode=0, partner=i))
            continue
        j = rng.randint(N)                                              # :175
        data2, t2 = images[j], labels[j]                                # :176
        if rng.random() < 0.5 and mixup_alpha > 0:                      # :185
            d, t, inf = mixup_ref(data1, data2, t1, t2, mixup_alpha, num_classes, rng)
        else:
            d, t, inf = cutmix_ref(data1, data2, t1, t2, cutmix_alpha, num_classes, rng)
        inf["partner"] = int(j)
        out[i], targets[i] = d, t
        info.append(inf)
    return out, targets, info


# ----------------------------------------------------------------------------------------------
# timm `Mixup(mode='batch')`.  PARITY UNPINNED: timm is an un-vendored dependency
# of the reference (requirements.txt:3 `timm>=0.9.2`; imported and constructed at
# previous/V2-convbase/dataset.py:10,119-129, never called) and is not installable here, so this
# restates the published algorithm of timm/data/mixup.py (`Mixup._params_per_batch`, `_mix_batch`,
# `rand_bbox`, `cutmix_bbox_and_lam`, `mixup_target`, `one_hot`) with torch ops on CPU tensors, in
# the upstream order of operations (in-place mul_/add_ on the batch, x.flip(0) as the partner).
def timm_batch_mixup_ref(x, target, *, mixup_alpha=1.0, cutmix_alpha=0.0, prob=1.0, switch_prob=0.5,
                         correct_lam=True, label_smoothing=0.1, num_classes=1000, rng=None):
    import torch
    rng = np.random if rng is None else rng
    assert len(x) % 2 == 0
    x = x.clone()
    lam, use_cutmix = 1.0, False
    if (mixup_alpha > 0 or cutmix_alpha > 0) and rng.rand() < prob:
        if mixup_alpha > 0.0 and cutmix_alpha > 0.0:
            use_cutmix = rng.rand() < switch_prob
            lam_mix = rng.beta(cutmix_alpha, cutmix_alpha) if use_cutmix else rng.beta(mixup_alpha, mixup_alpha)
        elif mixup_alpha > 0.0:
            lam_mix = rng.beta(mixup_alpha, mixup_alpha)
        else:
            use_cutmix = True
            lam_mix = rng.beta(cutmix_alpha, cutmix_alpha)
        lam = float(lam_mix)
    if lam != 1.0:
        if use_cutmix:
            img_h, img_w = x.shape[-2:]
            ratio = np.sqrt(1 - lam)
            cut_h, cut_w = int(img_h * ratio), int(img_w * ratio)
            cy = rng.randint(0, img_h)
            cx = rng.randint(0, img_w)
            yl, yh = np.clip(cy - cut_h // 2, 0, img_h), np.clip(cy + cut_h // 2, 0, img_h)
            xl, xh = np.clip(cx - cut_w // 2, 0, img_w), np.clip(cx + cut_w // 2, 0, img_w)
            if correct_lam:
                lam = 1.0 - (yh - yl) * (xh - xl) / float(img_h * img_w)
            x[:, :, yl:yh, xl:xh] = x.flip(0)[:, :, yl:yh, xl:xh]
        else:
            x_flipped = x.flip(0).mul_(1.0 - lam)
            x.mul_(lam).add_(x_flipped)
    off_value = label_smoothing / num_classes
    on_value = 1.0 - label_smoothing + off_value

    def one_hot(t):
        return torch.full((t.size(0), num_classes), off_value).scatter_(1, t.long().view(-1, 1), on_value)

    y = one_hot(target) * lam + one_hot(target.flip(0)) * (1.0 - lam)
    return x, y, float(lam)
