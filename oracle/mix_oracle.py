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
