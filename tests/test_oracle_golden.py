"""Pin the oracle: every restatement in oracle/ against the committed golden vectors that were
produced by the reference's own code (oracle/gen_golden.py).  CPU only."""
import math
import os

import numpy as np
import pytest
import torch

from oracle import block_oracle, mix_oracle
from image_classification_b200 import mix as mixmod


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


MIX_SEEDS = [0, 1, 2, 3, 4]


def _mix_case(g, seed):
    _, N, CH, H, W, NC = [int(v) for v in g[f"s{seed}_cfg"]]
    ma, ca, prob = [float(v) for v in g[f"s{seed}_alphas"]]
    gen = torch.Generator().manual_seed(seed)
    images = torch.randn(N, CH, H, W, generator=gen).numpy()
    labels = torch.randint(0, NC, (N,), generator=gen).numpy()
    return N, CH, H, W, NC, ma, ca, prob, images, labels


@pytest.mark.parametrize("seed", MIX_SEEDS)
def test_mix_oracle_bit_exact_vs_reference(golden_dir, seed):
    g = _load(golden_dir, "mix_reference.npz")
    N, CH, H, W, NC, ma, ca, prob, images, labels = _mix_case(g, seed)
    out, targets, _ = mix_oracle.mix_dataset_ref(images, labels, mixup_alpha=ma, cutmix_alpha=ca,
                                                 prob=prob, num_classes=NC,
                                                 rng=np.random.RandomState(seed))
    assert np.array_equal(targets, g[f"s{seed}_targets"])
    if f"s{seed}_mixed" in g:
        assert np.array_equal(out, g[f"s{seed}_mixed"])
    else:
        assert np.array_equal(out[:, :, H // 2, :], g[f"s{seed}_mixed_row"])
        assert np.array_equal(out.astype(np.float64).sum(axis=(1, 2, 3)), g[f"s{seed}_mixed_sum"])


@pytest.mark.parametrize("seed", MIX_SEEDS)
def test_host_planner_matches_oracle_rng_order(golden_dir, seed):
    """The product's planner draws the same integers (mode, partner, box) and the same float32
    lambdas as the oracle under the same RandomState — bit-exact."""
    g = _load(golden_dir, "mix_reference.npz")
    N, CH, H, W, NC, ma, ca, prob, images, labels = _mix_case(g, seed)
    _, targets, info = mix_oracle.mix_dataset_ref(images, labels, mixup_alpha=ma, cutmix_alpha=ca,
                                                  prob=prob, num_classes=NC,
                                                  rng=np.random.RandomState(seed))
    plan = mixmod.plan_mix_batch(N, H, W, mixup_alpha=ma, cutmix_alpha=ca, prob=prob,
                                 rng=np.random.RandomState(seed))
    for i, inf in enumerate(info):
        assert plan["mode"][i] == inf["mode"]
        assert plan["partner"][i] == inf["partner"]
        if inf["mode"] == 2:
            assert (plan["bbx1"][i], plan["bby1"][i], plan["bbx2"][i], plan["bby2"][i]) == inf["box"]
        if inf["mode"] != 0:
            assert plan["lam"][i] == np.float32(inf["lam"])
            assert plan["one_minus_lam"][i] == np.float32(1 - inf["lam"])
    # targets rebuilt from the plan alone reproduce the reference rows bit for bit
    rebuilt = np.zeros((N, NC), dtype=np.float32)
    for i in range(N):
        if plan["mode"][i] == 0:
            rebuilt[i, labels[i]] = 1.0
        else:
            a = np.zeros(NC, np.float32); a[labels[i]] = plan["lam"][i]
            b = np.zeros(NC, np.float32); b[labels[plan["partner"][i]]] = plan["one_minus_lam"][i]
            rebuilt[i] = a + b
    assert np.array_equal(rebuilt, g[f"s{seed}_targets"])


def test_global_numpy_rng_is_the_default(golden_dir):
    """Reference code calls np.random.* directly; rng=None must consume that same stream."""
    np.random.seed(7)
    a = mixmod.plan_mix_batch(16, 8, 8, mixup_alpha=0.2, cutmix_alpha=1.0, prob=0.5)
    b = mixmod.plan_mix_batch(16, 8, 8, mixup_alpha=0.2, cutmix_alpha=1.0, prob=0.5,
                              rng=np.random.RandomState(7))
    assert a.tobytes() == b.tobytes()


def test_ema_oracle_bit_exact_vs_reference(golden_dir):
    g = _load(golden_dir, "ema_reference.npz")
    decay = float(g["decay"])
    names = sorted(k[len("step0_param_"):] for k in g.files if k.startswith("step0_param_"))
    assert names
    for name in names:
        shadow = torch.from_numpy(g[f"init_shadow_{name}"])
        for step in range(3):
            p = torch.from_numpy(g[f"step{step}_param_{name}"])
            shadow = block_oracle.ema_update_ref(shadow, p, decay)
            assert torch.equal(shadow, torch.from_numpy(g[f"step{step}_shadow_{name}"])), name


def test_softce_oracle_vs_torch_criterion(golden_dir):
    g = _load(golden_dir, "softce_torch.npz")
    for c in range(4):
        z = torch.from_numpy(g[f"c{c}_z"]).requires_grad_()
        t = torch.from_numpy(g[f"c{c}_t"])
        eps = float(g[f"c{c}_eps"])
        loss = block_oracle.soft_ce_ref(z, t, eps)
        loss.backward()
        np.testing.assert_allclose(loss.detach().numpy(), g[f"c{c}_loss"], rtol=1e-6, atol=1e-7)
        np.testing.assert_allclose(z.grad.numpy(), g[f"c{c}_dz"], rtol=1e-5, atol=1e-8)


def test_schedule_mirror_vs_reference(golden_dir):
    from image_classification_b200.train_step import cosine_schedule_with_warmup_lambda
    g = _load(golden_dir, "schedule_reference.npz")
    lam = cosine_schedule_with_warmup_lambda(5, 40, 1e-6)
    lrs = np.array([1e-4 * lam(s) for s in range(45)])
    np.testing.assert_allclose(lrs, g["lrs"], rtol=1e-12, atol=0)


@pytest.mark.parametrize("case", [0, 1])
def test_block_oracle_vs_torchvision_cnblock(golden_dir, case):
    """fp64 forward + all gradients of the restated block equal torchvision's CNBlock."""
    g = _load(golden_dir, "block_torchvision.npz")
    P = lambda k: torch.from_numpy(g[f"b{case}_p_{k}"]).requires_grad_()
    x = torch.from_numpy(g[f"b{case}_x"]).requires_grad_()
    params = dict(conv_w=P("block.0.weight"), conv_b=P("block.0.bias"), norm_w=P("block.2.weight"),
                  norm_b=P("block.2.bias"), fc1_w=P("block.3.weight"), fc1_b=P("block.3.bias"),
                  fc2_w=P("block.5.weight"), fc2_b=P("block.5.bias"))
    ls = P("layer_scale")
    y = block_oracle.block_forward_ref(x, gamma=ls.reshape(-1), **params)
    np.testing.assert_allclose(y.detach().numpy(), g[f"b{case}_y"], rtol=1e-10, atol=1e-12)
    y.backward(torch.from_numpy(g[f"b{case}_gy"]))
    np.testing.assert_allclose(x.grad.numpy(), g[f"b{case}_dx"], rtol=1e-9, atol=1e-12)
    names = {"conv_w": "block.0.weight", "conv_b": "block.0.bias", "norm_w": "block.2.weight",
             "norm_b": "block.2.bias", "fc1_w": "block.3.weight", "fc1_b": "block.3.bias",
             "fc2_w": "block.5.weight", "fc2_b": "block.5.bias"}
    for k, tvk in names.items():
        np.testing.assert_allclose(params[k].grad.numpy(), g[f"b{case}_g_{tvk}"], rtol=1e-8,
                                   atol=1e-12, err_msg=k)
    np.testing.assert_allclose(ls.grad.numpy(), g[f"b{case}_g_layer_scale"], rtol=1e-8, atol=1e-12)


def test_dwconv_ln_ref_consistent_with_block():
    """dwconv_ln_ref (no affine) + explicit affine == the first half of the block."""
    torch.manual_seed(0)
    B, C, H, W = 2, 16, 9, 11
    x = torch.randn(B, H, W, C, dtype=torch.float64)
    w = torch.randn(C, 1, 7, 7, dtype=torch.float64) * 0.2
    b = torch.randn(C, dtype=torch.float64)
    xhat, mean, rstd = block_oracle.dwconv_ln_ref(x, w, b)
    y = torch.nn.functional.conv2d(x.permute(0, 3, 1, 2), w, b, padding=3, groups=C).permute(0, 2, 3, 1)
    ref = torch.nn.functional.layer_norm(y, (C,), None, None, 1e-6)
    assert torch.allclose(xhat, ref, atol=1e-10)
    assert math.isclose(float(mean[0]), float(y[0, 0, 0].mean()), rel_tol=1e-10)
