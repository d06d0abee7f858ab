"""Generate tests/golden/*.npz by running the REFERENCE'S OWN code (from /root/reference) and
torch's own criterion on seeded inputs.  Run in this container:  python -m oracle.gen_golden

TEST INFRASTRUCTURE.  The fixtures pin oracle/ (tests/test_oracle_golden.py) and travel to the
GPU box, where /root/reference does not exist, so the `-m gpu` parity tests can compare the CUDA
path with reference outputs directly.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from .ref_loader import load_reference

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


class _MemDataset(torch.utils.data.Dataset):
    def __init__(self, images, labels):
        self.images, self.labels = images, labels

    def __len__(self):
        return self.images.shape[0]

    def __getitem__(self, i):
        return self.images[i], self.labels[i]


def gen_mix(ds_mod):
    """MixUpCutMixDataset over in-memory datasets, iterated in order under np.random.seed."""
    cases = [  # (seed, N, CH, H, W, num_classes, mixup_alpha, cutmix_alpha, prob)
        (0, 16, 3, 12, 20, 44, 0.2, 1.0, 0.5),      # train_advanced_v4.py:593-599 settings
        (1, 32, 3, 32, 32, 44, 0.2, 1.0, 0.5),
        (2, 8, 3, 8, 16, 10, 1.0, 1.0, 1.0),        # always mixed
        (3, 8, 1, 16, 8, 5, 0.0, 1.0, 1.0),         # mixup disabled -> cutmix only (:185)
        (4, 24, 3, 224, 224, 44, 0.2, 1.0, 0.5),    # real image size (plan only is stored)
    ]
    out = {}
    for (seed, N, CH, H, W, NC, ma, ca, prob) in cases:
        g = torch.Generator().manual_seed(seed)
        images = torch.randn(N, CH, H, W, generator=g)
        labels = torch.randint(0, NC, (N,), generator=g)
        m = ds_mod.MixUpCutMixDataset(_MemDataset(images, labels), mixup_alpha=ma, cutmix_alpha=ca,
                                      prob=prob, num_classes=NC)
        np.random.seed(seed)
        items = [m[i] for i in range(N)]
        mixed = torch.stack([it[0] for it in items]).numpy()
        targets = torch.stack([it[1] for it in items]).numpy()
        key = f"s{seed}"
        out[f"{key}_cfg"] = np.array([seed, N, CH, H, W, NC], dtype=np.int64)
        out[f"{key}_alphas"] = np.array([ma, ca, prob], dtype=np.float64)
        out[f"{key}_targets"] = targets
        if H * W <= 1024:
            out[f"{key}_mixed"] = mixed
        else:   # keep the fixture small: a checksum of the pixels instead of the pixels
            out[f"{key}_mixed_sum"] = mixed.astype(np.float64).sum(axis=(1, 2, 3))
            out[f"{key}_mixed_row"] = mixed[:, :, H // 2, :]
    np.savez_compressed(os.path.join(GOLDEN, "mix_reference.npz"), **out)


def gen_ema(tr_mod):
    """EMA.register/update x3 on a tiny model (train_advanced_v4.py:62-80)."""
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.Conv2d(3, 4, 3), torch.nn.LayerNorm(9))
    ema = tr_mod.EMA(model, 0.9997)
    out = {"decay": np.array(0.9997)}
    for name, s0 in ema.shadow.items():
        out[f"init_shadow_{name}"] = s0.numpy().copy()
    for step in range(3):
        with torch.no_grad():
            for i, p in enumerate(model.parameters()):
                g = torch.Generator().manual_seed(100 * step + i)
                p.add_(torch.randn(p.shape, generator=g))
        for name, p in model.named_parameters():
            out[f"step{step}_param_{name}"] = p.detach().numpy().copy()
        ema.update()
        for name, s in ema.shadow.items():
            out[f"step{step}_shadow_{name}"] = s.numpy().copy()
    np.savez_compressed(os.path.join(GOLDEN, "ema_reference.npz"), **out)


def gen_schedule(tr_mod):
    """get_cosine_schedule_with_warmup (train_advanced_v4.py:183-191)."""
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.AdamW([p], lr=1e-4)
    sch = tr_mod.get_cosine_schedule_with_warmup(opt, num_warmup_steps=5, num_training_steps=40,
                                                 min_lr=1e-6)
    lrs = []
    for _ in range(45):
        lrs.append(opt.param_groups[0]["lr"])
        opt.step()
        sch.step()
    np.savez_compressed(os.path.join(GOLDEN, "schedule_reference.npz"), lrs=np.array(lrs))


def gen_softce():
    """nn.CrossEntropyLoss(label_smoothing=0.1) on probability targets
    (train_advanced_v4.py:626, config.py:31) — torch itself is the reference implementation."""
    out = {}
    for seed, (B, NC, eps) in enumerate([(8, 44, 0.1), (128, 44, 0.1), (5, 1000, 0.0), (3, 7, 0.3)]):
        g = torch.Generator().manual_seed(seed)
        z = (torch.randn(B, NC, generator=g) * 3).requires_grad_()
        t = torch.softmax(torch.randn(B, NC, generator=g) * 2, dim=1)
        if seed == 0:   # mixup-like two-hot rows
            t = torch.zeros(B, NC)
            for b in range(B):
                lam = 0.1 + 0.1 * b
                t[b, b % NC] += lam
                t[b, (3 * b + 1) % NC] += 1 - lam
        loss = torch.nn.CrossEntropyLoss(label_smoothing=eps)(z, t)
        loss.backward()
        out[f"c{seed}_z"] = z.detach().numpy()
        out[f"c{seed}_t"] = t.numpy()
        out[f"c{seed}_eps"] = np.array(eps)
        out[f"c{seed}_loss"] = loss.detach().numpy()
        out[f"c{seed}_dz"] = z.grad.numpy()
    np.savez_compressed(os.path.join(GOLDEN, "softce_torch.npz"), **out)


def gen_block_torchvision():
    """torchvision CNBlock forward/backward (same architecture as timm's block) in float64."""
    from torchvision.models.convnext import CNBlock
    out = {}
    for case, (B, C, H, W) in enumerate([(2, 64, 9, 7), (1, 96, 14, 14)]):
        torch.manual_seed(case)
        blk = CNBlock(C, layer_scale=1.0, stochastic_depth_prob=0.0).double()
        with torch.no_grad():
            for p in blk.parameters():
                p.copy_(torch.randn_like(p) * 0.2)
            blk.layer_scale.copy_(torch.rand_like(blk.layer_scale) * 0.9 + 0.1)
        x = torch.randn(B, C, H, W, dtype=torch.float64, requires_grad=True)
        y = blk(x)
        gy = torch.randn_like(y)
        y.backward(gy)
        sd = {k: v.detach().numpy() for k, v in blk.state_dict().items()}
        out[f"b{case}_x"] = x.detach().numpy()
        out[f"b{case}_gy"] = gy.numpy()
        out[f"b{case}_y"] = y.detach().numpy()
        out[f"b{case}_dx"] = x.grad.numpy()
        for k, v in sd.items():
            out[f"b{case}_p_{k}"] = v
        for k, p in blk.named_parameters():
            out[f"b{case}_g_{k}"] = p.grad.numpy()
    np.savez_compressed(os.path.join(GOLDEN, "block_torchvision.npz"), **out)


def loop_fixture_inputs(n_micro=6, batch=4, num_classes=10, seed=11):
    """Seeded inputs of the loop fixture: [(images [b,3,8,8], soft labels [b,NC])] * n_micro."""
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n_micro):
        x = torch.randn(batch, 3, 8, 8, generator=g)
        t = torch.softmax(torch.randn(batch, num_classes, generator=g) * 2, dim=1)
        out.append((x, t))
    return out


def gen_loop(tr_mod):
    """The reference's OWN `train_one_epoch` (train_advanced_v4.py:193-281) with its `EMA`
    (:62-94) and `get_cosine_schedule_with_warmup` (:183-191), 6 (and 5: odd tail) micro-batches of
    soft labels on CPU over a tiny model that returns a list (hits :219-220 / :241-242), once with
    scaler=None (plain branch, :238-258) and once with a disabled GradScaler (AMP branch, :215-237:
    un-divided loss at :223).  Parameters after every micro-batch, final EMA shadows, the LR trace
    and the returned mean loss are stored."""
    import types
    from .loop_oracle import TinyListModel
    out = {}
    cfg = types.SimpleNamespace(num_classes=10, gradient_accumulation_steps=2, gradient_clip_val=1.0)
    torch.manual_seed(3)
    init = TinyListModel().state_dict()
    for k, v in init.items():
        out[f"init_{k}"] = v.numpy().copy()
    for case, (amp, n_micro) in {"plain": (False, 6), "amp": (True, 6), "plain_odd": (False, 5)}.items():
        model = TinyListModel()
        model.load_state_dict(init)
        opt = torch.optim.AdamW(model.parameters(), lr=5e-3, weight_decay=1e-2)
        sched = tr_mod.get_cosine_schedule_with_warmup(opt, 1, 3, min_lr=1e-6)
        ema = tr_mod.EMA(model, 0.9)
        crit = torch.nn.CrossEntropyLoss(label_smoothing=0.1)
        scaler = torch.amp.GradScaler("cpu", enabled=False) if amp else None
        batches = loop_fixture_inputs(n_micro)
        snaps, lrs = [], []

        def snap():
            snaps.append({k: v.detach().numpy().copy() for k, v in model.named_parameters()})
            lrs.append(opt.param_groups[0]["lr"])

        def loader():
            for i, b in enumerate(batches):
                if i > 0:
                    snap()              # state after micro-batch i-1
                yield b

        mean_loss, _ = tr_mod.train_one_epoch(model, loader(), crit, opt, sched, "cpu", cfg,
                                              scaler=scaler, ema=ema)
        snap()
        assert len(snaps) == n_micro
        for i, sn in enumerate(snaps):
            for k, v in sn.items():
                out[f"{case}_m{i}_{k}"] = v
        out[f"{case}_lrs"] = np.array(lrs)
        out[f"{case}_mean_loss"] = np.array(mean_loss)
        out[f"{case}_n_micro"] = np.array(n_micro)
        for k, v in ema.shadow.items():
            out[f"{case}_shadow_{k}"] = v.numpy().copy()
        for k, p in model.named_parameters():      # leftover gradients of the odd tail (:225-231)
            if p.grad is not None:
                out[f"{case}_grad_{k}"] = p.grad.numpy().copy()
    np.savez_compressed(os.path.join(GOLDEN, "loop_reference.npz"), **out)


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    ds_mod, tr_mod = load_reference()
    gen_mix(ds_mod)
    gen_ema(tr_mod)
    gen_schedule(tr_mod)
    gen_loop(tr_mod)
    gen_softce()
    gen_block_torchvision()
    for f in sorted(os.listdir(GOLDEN)):
        print(f, os.path.getsize(os.path.join(GOLDEN, f)))


if __name__ == "__main__":
    main()
