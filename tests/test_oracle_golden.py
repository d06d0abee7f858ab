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


# ------------------------------------------------------------------ the step loop (row a6)
LOOP_CASES = [("plain", False), ("amp", True), ("plain_odd", False)]


def _loop_fixture(golden_dir):
    from oracle.gen_golden import loop_fixture_inputs
    from oracle.loop_oracle import TinyListModel
    g = _load(golden_dir, "loop_reference.npz")
    model = TinyListModel()
    model.load_state_dict({k[len("init_"):]: torch.from_numpy(g[k]) for k in g.files
                           if k.startswith("init_")})
    return g, model, loop_fixture_inputs


def _check_loop_case(g, case, model, ema_shadow, lrs, mean_loss, snaps, tol):
    n = int(g[f"{case}_n_micro"])
    assert len(snaps) == n and len(lrs) == n
    for i, sn in enumerate(snaps):
        for k, v in sn.items():
            ref = g[f"{case}_m{i}_{k}"]
            assert np.allclose(v, ref, rtol=tol, atol=tol), (case, i, k, np.abs(v - ref).max())
    assert np.allclose(np.array(lrs), g[f"{case}_lrs"], rtol=1e-12, atol=0)
    assert abs(mean_loss - float(g[f"{case}_mean_loss"])) < tol * 10
    for k, s in ema_shadow.items():
        ref = g[f"{case}_shadow_{k}"]
        assert np.allclose(s, ref, rtol=tol, atol=tol), (case, "shadow", k)
    # the odd tail leaves its gradients in .grad (train_advanced_v4.py:225-231 never ran)
    for k, p in model.named_parameters():
        key = f"{case}_grad_{k}"
        if key in g.files:
            assert p.grad is not None and np.allclose(p.grad.detach().cpu().numpy(), g[key],
                                                      rtol=tol, atol=tol), (case, "grad", k)
        else:
            assert p.grad is None or float(p.grad.abs().sum()) == 0.0, (case, "grad", k)


@pytest.mark.parametrize("case,amp", LOOP_CASES)
def test_loop_oracle_reproduces_reference_train_one_epoch(golden_dir, case, amp):
    """oracle/loop_oracle.reference_loop vs the reference's OWN train_one_epoch + EMA + schedule
    (fixture: oracle/gen_golden.py::gen_loop): parameters after every micro-batch, EMA shadows,
    LR trace, returned mean loss — both branches and the odd-tail case."""
    from oracle import loop_oracle
    g, model, inputs = _loop_fixture(golden_dir)
    n = int(g[f"{case}_n_micro"])
    opt = torch.optim.AdamW(model.parameters(), lr=5e-3, weight_decay=1e-2)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, loop_oracle.cosine_warmup_lambda(1, 3, 1e-6))
    ema = loop_oracle.EMARef(model, 0.9)
    snaps, lrs = [], []

    def on_micro(i):
        snaps.append({k: v.detach().numpy().copy() for k, v in model.named_parameters()})
        lrs.append(opt.param_groups[0]["lr"])

    mean_loss = loop_oracle.reference_loop(model, inputs(n), torch.nn.CrossEntropyLoss(label_smoothing=0.1),
                                           opt, sched, ema, accum=2, clip=1.0, amp_branch=amp,
                                           on_micro=on_micro)
    _check_loop_case(g, case, model, {k: v.numpy() for k, v in ema.shadow.items()}, lrs, mean_loss,
                     snaps, 1e-6)


@pytest.mark.parametrize("case,amp", LOOP_CASES)
def test_train_step_reproduces_reference_train_one_epoch_cpu(golden_dir, case, amp):
    """The PRODUCT's `TrainStep.micro_step` (torch optimizer calls; the fused CUDA step is covered
    by tests/test_gpu_loop.py) against the same fixture, on CPU tensors."""
    import warnings
    from image_classification_b200 import train_step
    from oracle import loop_oracle
    g, model, inputs = _loop_fixture(golden_dir)
    n = int(g[f"{case}_n_micro"])
    opt = torch.optim.AdamW(model.parameters(), lr=5e-3, weight_decay=1e-2)
    sched = train_step.get_cosine_schedule_with_warmup(opt, 1, 3, 1e-6)
    ema = loop_oracle.EMARef(model, 0.9)
    scaler = torch.amp.GradScaler("cpu", enabled=False) if amp else None
    step = train_step.TrainStep(model, torch.nn.CrossEntropyLoss(label_smoothing=0.1), opt, sched,
                                scaler, ema, train_step.StepConfig(2, 1.0, 10),
                                use_fused_optim=False)
    snaps, lrs, s, cnt = [], [], 0.0, 0
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")            # autocast('cuda') on a CPU-only machine
        for i, (x, t) in enumerate(inputs(n)):
            loss = step.micro_step(x, t, i)
            s += float(loss) * x.size(0)
            cnt += x.size(0)
            snaps.append({k: v.detach().numpy().copy() for k, v in model.named_parameters()})
            lrs.append(opt.param_groups[0]["lr"])
    step.close()
    _check_loop_case(g, case, model, {k: v.numpy() for k, v in ema.shadow.items()}, lrs, s / cnt,
                     snaps, 1e-6)
