"""GPU parity (through the C ABI) of the glue kernels: mix, soft-CE, EMA, layout, weight folding.
Bit-exact where the reference is integer/deterministic fp32 (mix, EMA); 1e-4 relative in fp32."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import block_oracle, mix_oracle  # noqa: E402  (checker only)


def _rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a B200"
    return torch.device("cuda:0")


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


# ------------------------------------------------------------------------------------- mix
@pytest.mark.parametrize("seed", [0, 1, 2, 3, 4])
def test_mix_batch_bit_exact_vs_reference_golden(cnx_lib, golden_dir, seed):
    from image_classification_b200 import mix
    g = _load(golden_dir, "mix_reference.npz")
    _, N, CH, H, W, NC = [int(v) for v in g[f"s{seed}_cfg"]]
    ma, ca, prob = [float(v) for v in g[f"s{seed}_alphas"]]
    gen = torch.Generator().manual_seed(seed)
    images = torch.randn(N, CH, H, W, generator=gen)
    labels = torch.randint(0, NC, (N,), generator=gen)
    plan = mix.plan_mix_batch(N, H, W, mixup_alpha=ma, cutmix_alpha=ca, prob=prob,
                              rng=np.random.RandomState(seed))
    out, targets = mix.mix_batch(images.to(_dev()), labels.to(_dev()), plan, NC)
    out, targets = out.cpu().numpy(), targets.cpu().numpy()
    assert np.array_equal(targets, g[f"s{seed}_targets"])
    if f"s{seed}_mixed" in g:
        assert np.array_equal(out, g[f"s{seed}_mixed"])
    else:
        assert np.array_equal(out[:, :, H // 2, :], g[f"s{seed}_mixed_row"])
        assert np.array_equal(out.astype(np.float64).sum(axis=(1, 2, 3)), g[f"s{seed}_mixed_sum"])


def test_mix_batch_full_size_vs_oracle(cnx_lib):
    """BASELINE config 2 shape (B=128, 3x224x224), reference settings, bit-exact vs the oracle."""
    from image_classification_b200 import mix
    N, CH, H, W, NC = 128, 3, 224, 224, 44
    gen = torch.Generator().manual_seed(0)
    images = torch.randn(N, CH, H, W, generator=gen)
    labels = torch.randint(0, NC, (N,), generator=gen)
    ref, ref_t, _ = mix_oracle.mix_dataset_ref(images.numpy(), labels.numpy(), mixup_alpha=0.2,
                                               cutmix_alpha=1.0, prob=0.5, num_classes=NC,
                                               rng=np.random.RandomState(0))
    mixer = mix.BatchMixUpCutMix(mixup_alpha=0.2, cutmix_alpha=1.0, prob=0.5, num_classes=NC)
    out, targets = mixer(images.to(_dev()), labels.to(_dev()), rng=np.random.RandomState(0))
    assert np.array_equal(out.cpu().numpy(), ref)
    assert np.array_equal(targets.cpu().numpy(), ref_t)
    # size-independent properties: rows sum to 1 (within 1 ulp), passthrough rows are one-hot
    assert torch.all((targets.sum(1) - 1).abs() < 1e-6)


def test_mix_rejects_bad_arguments(cnx_lib):
    from image_classification_b200 import _lib, mix
    x = torch.zeros(2, 3, 4, 6, device=_dev())   # W not a multiple of 4
    plan = mix.plan_mix_batch(2, 4, 6, rng=np.random.RandomState(0))
    with pytest.raises(_lib.CnxError):
        mix.mix_batch(x, torch.zeros(2, dtype=torch.long, device=_dev()), plan, 5)
    with pytest.raises(_lib.CnxError):
        mix.mix_batch(torch.zeros(2, 3, 4, 8), torch.zeros(2, dtype=torch.long), plan, 5)  # CPU


# ---------------------------------------------------------------------------------- soft CE
@pytest.mark.parametrize("case", [0, 1, 2, 3])
def test_softce_vs_torch_golden(cnx_lib, golden_dir, case):
    from image_classification_b200.loss import SoftTargetCrossEntropyB200
    g = _load(golden_dir, "softce_torch.npz")
    z = torch.from_numpy(g[f"c{case}_z"]).to(_dev()).requires_grad_()
    t = torch.from_numpy(g[f"c{case}_t"]).to(_dev())
    crit = SoftTargetCrossEntropyB200(label_smoothing=float(g[f"c{case}_eps"]))
    loss = crit(z, t)
    (loss * 3.0).backward()      # exercises the upstream-gradient scaling (GradScaler)
    np.testing.assert_allclose(loss.item(), g[f"c{case}_loss"], rtol=1e-4)          # fp32: 1e-4
    ref_dz = g[f"c{case}_dz"] * 3.0
    err = np.abs(z.grad.cpu().numpy() - ref_dz).max() / np.abs(ref_dz).max()
    assert err < 1e-4, err


def test_softce_hard_labels_and_bf16_logits(cnx_lib):
    from image_classification_b200.loss import SoftTargetCrossEntropyB200
    torch.manual_seed(0)
    z = torch.randn(16, 44, device=_dev()).bfloat16().requires_grad_()
    y = torch.randint(0, 44, (16,), device=_dev())
    loss = SoftTargetCrossEntropyB200(0.1)(z, y)
    ref = torch.nn.functional.cross_entropy(z.float(), y, label_smoothing=0.1)
    assert abs(loss.item() - ref.item()) < 1e-4 * abs(ref.item())
    loss.backward()
    assert z.grad.dtype == torch.bfloat16


# -------------------------------------------------------------------------------------- EMA
def test_ema_bit_exact_vs_reference_golden(cnx_lib, golden_dir):
    from image_classification_b200.ema import EMAB200
    g = _load(golden_dir, "ema_reference.npz")
    model = torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.Conv2d(3, 4, 3),
                                torch.nn.LayerNorm(9)).to(_dev())
    with torch.no_grad():
        for name, p in model.named_parameters():
            p.copy_(torch.from_numpy(g[f"init_shadow_{name}"]))
    ema = EMAB200(model, float(g["decay"]))
    for step in range(3):
        with torch.no_grad():
            for name, p in model.named_parameters():
                p.copy_(torch.from_numpy(g[f"step{step}_param_{name}"]))
        ema.update()
        for name, s in ema.shadow.items():
            assert np.array_equal(s.cpu().numpy(), g[f"step{step}_shadow_{name}"]), (step, name)


def test_ema_many_ragged_tensors_and_swap(cnx_lib):
    from image_classification_b200.ema import EMAB200
    torch.manual_seed(1)
    sizes = [1, 3, 4, 5, 8191, 8192, 8193, 3 * 8192 + 7, 100_003, 44, 1024 * 49]
    model = torch.nn.ParameterList([torch.nn.Parameter(torch.randn(n, device=_dev())) for n in sizes])
    frozen = torch.nn.Parameter(torch.randn(10, device=_dev()), requires_grad=False)
    model.append(frozen)
    ema = EMAB200(model, 0.9997)
    assert len(ema.shadow) == len(sizes)                  # requires_grad only (:72)
    ref = {k: v.clone() for k, v in ema.shadow.items()}
    for _ in range(2):
        with torch.no_grad():
            for p in model:
                if p.requires_grad:
                    p.add_(torch.randn_like(p))
        ema.update()
        for (k, s), p in zip(ref.items(), [p for p in model if p.requires_grad]):
            ref[k] = block_oracle.ema_update_ref(s, p.data, 0.9997)
    for k in ref:
        assert torch.equal(ema.shadow[k], ref[k]), k
    before = [p.data.clone() for p in model]
    ema.apply_shadow()
    assert all(torch.equal(p.data, ema.shadow[n]) for n, p in model.named_parameters() if p.requires_grad)
    ema.restore()
    assert all(torch.equal(p.data, b) for p, b in zip(model, before))


# ----------------------------------------------------------------------------------- layout
def test_layout_round_trip(cnx_lib):
    from image_classification_b200 import ops
    torch.manual_seed(0)
    for shape in [(2, 96, 7, 5), (1, 128, 14, 14), (3, 40, 9, 33)]:
        x = torch.randn(*shape, device=_dev())
        nhwc = ops.nchw_f32_to_nhwc_bf16(x)
        assert torch.equal(nhwc, x.permute(0, 2, 3, 1).contiguous().bfloat16())
        back = ops.nhwc_bf16_to_nchw_f32(nhwc)
        assert torch.equal(back, x.bfloat16().float())


# ----------------------------------------------------------------- weight folding / adjoint
def _rand_block_params(C, dev, gamma_scale=1.0):
    torch.manual_seed(C)
    f = lambda *s, std=0.2: (torch.randn(*s, device=dev) * std)
    return dict(conv_w=f(C, 1, 7, 7), conv_b=f(C), norm_w=1 + f(C), norm_b=f(C),
                fc1_w=f(4 * C, C, std=0.1), fc1_b=f(4 * C), fc2_w=f(C, 4 * C, std=0.1),
                fc2_b=f(C), gamma=(torch.rand(C, device=dev) * 0.9 + 0.1) * gamma_scale)


@pytest.mark.parametrize("C", [64, 96, 128, 512])
def test_prep_block_weights(cnx_lib, C):
    from image_classification_b200 import ops
    p = _rand_block_params(C, _dev())
    fw = ops.prep_block_weights(p["conv_w"], p["norm_w"], p["norm_b"], p["fc1_w"], p["fc1_b"],
                                p["fc2_w"], p["fc2_b"], p["gamma"])
    w1 = (p["fc1_w"] * p["norm_w"][None, :]).bfloat16()
    w2 = (p["gamma"][:, None] * p["fc2_w"]).bfloat16()
    assert torch.equal(fw.w1f, w1) and torch.equal(fw.w1fT, w1.t().contiguous())
    assert torch.equal(fw.w2f, w2) and torch.equal(fw.w2fT, w2.t().contiguous())
    assert torch.equal(fw.wt_dw, p["conv_w"].reshape(C, 49).t().contiguous())
    torch.testing.assert_close(fw.b1f, p["fc1_b"] + p["fc1_w"] @ p["norm_b"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(fw.b2f, p["gamma"] * p["fc2_b"], rtol=1e-6, atol=0)


@pytest.mark.parametrize("C", [64, 96, 256])
def test_unfold_block_grads_is_adjoint_of_prep(cnx_lib, C):
    from image_classification_b200 import ops
    dev = _dev()
    p = {k: v.double().requires_grad_() for k, v in _rand_block_params(C, dev).items()}
    w1f = p["fc1_w"] * p["norm_w"][None, :]
    b1f = p["fc1_b"] + p["fc1_w"] @ p["norm_b"]
    w2f = p["gamma"][:, None] * p["fc2_w"]
    b2f = p["gamma"] * p["fc2_b"]
    wt = p["conv_w"].reshape(C, 49).t()
    torch.manual_seed(1)
    up = [torch.randn_like(t) for t in (w1f, b1f, w2f, b2f, wt)]
    torch.autograd.backward([w1f, b1f, w2f, b2f, wt], up)
    f32 = lambda t: t.detach().float().contiguous()
    d_dw, d_nw, d_nb, d_f1, d_f2, d_f2b, d_g = ops.unfold_block_grads(
        f32(up[0]), f32(up[1]), f32(up[2]), f32(up[3]), f32(up[4]), f32(p["norm_w"]),
        f32(p["norm_b"]), f32(p["fc1_w"]), f32(p["fc2_w"]), f32(p["fc2_b"]), f32(p["gamma"]))
    for got, key in ((d_dw, "conv_w"), (d_nw, "norm_w"), (d_nb, "norm_b"), (d_f1, "fc1_w"),
                     (d_f2, "fc2_w"), (d_f2b, "fc2_b"), (d_g, "gamma")):
        ref = p[key].grad.float()
        err = (got - ref).abs().max() / ref.abs().max()
        assert err < 1e-4, (key, float(err))


# ---------------------------------------------------------------- LayerNorm2d (stem / downsample)
@pytest.mark.parametrize("shape", [(301, 64), (1000, 128), (77, 96), (513, 256), (260, 512),
                                   (130, 1024), (5, 8)])
def test_layernorm_affine_fwd_bwd_vs_torch_fp64(cnx_lib, shape):
    """cnx_layernorm_fwd/bwd vs F.layer_norm in fp64 on the same bf16-representable inputs."""
    import torch.nn.functional as F
    from image_classification_b200 import ops
    M, C = shape
    dev = torch.device("cuda:0")
    torch.manual_seed(M + C)
    x = (torch.randn(M, C, device=dev) * 1.7 + 0.4).bfloat16()
    g = torch.randn(M, C, device=dev).bfloat16()
    w = (1 + 0.3 * torch.randn(C, device=dev))
    b = 0.2 * torch.randn(C, device=dev)
    xr = x.double().requires_grad_()
    wr, br = w.double().requires_grad_(), b.double().requires_grad_()
    yr = F.layer_norm(xr, (C,), wr, br, 1e-6)
    yr.backward(g.double())
    y, mean, rstd = ops.layernorm_fwd(x, w, b, 1e-6)
    dx, dw, db = ops.layernorm_bwd(g, x, w, mean, rstd)
    rel = lambda a, r: float((a.double() - r).abs().max() / r.abs().max().clamp_min(1e-30))
    assert rel(mean, x.double().mean(1)) < 1e-4 or (mean.double() - x.double().mean(1)).abs().max() < 1e-5
    assert rel(rstd, torch.rsqrt(x.double().var(1, unbiased=False) + 1e-6)) < 1e-4
    assert rel(y, yr.detach()) < 2e-2
    assert (y.double() - yr.detach()).abs().max() < 2.0 ** -8 * yr.abs().max() + 1e-3
    assert rel(dx, xr.grad) < 2e-2
    assert rel(dw, wr.grad) < 1e-3 and rel(db, br.grad) < 1e-3


def test_layernorm2d_module_uses_library_and_matches_torch(cnx_lib):
    from image_classification_b200 import _lib
    from image_classification_b200.convnext import LayerNorm2d
    dev = torch.device("cuda:0")
    torch.manual_seed(3)
    ln = LayerNorm2d(128, eps=1e-6).to(dev)
    with torch.no_grad():
        ln.weight.normal_(1, 0.2)
        ln.bias.normal_(0, 0.2)
    x = torch.randn(2, 128, 9, 7, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    xa = x.clone().requires_grad_()
    n0 = _lib.launch_count
    y = ln(xa)
    y.float().square().sum().backward()
    assert _lib.launch_count - n0 == 2                       # fwd + bwd kernels of libcnx_b200
    xb = x.double().requires_grad_()
    yb = torch.nn.functional.layer_norm(xb.permute(0, 2, 3, 1), (128,), ln.weight.double(),
                                        ln.bias.double(), 1e-6).permute(0, 3, 1, 2)
    yb.square().sum().backward()
    rel = lambda a, r: float((a.double() - r).abs().max() / r.abs().max())
    assert y.dtype == torch.bfloat16 and rel(y, yb.detach()) < 2e-2
    assert rel(xa.grad, xb.grad) < 2e-2


# ------------------------------------------------------------------- fused optimizer step
class _RefEMA:
    """The reference's EMA.update expression (train_advanced_v4.py:75-80) in plain torch."""

    def __init__(self, model, decay):
        self.model, self.decay = model, decay
        self.shadow = {n: p.data.clone() for n, p in model.named_parameters() if p.requires_grad}

    def update(self):
        for n, p in self.model.named_parameters():
            if p.requires_grad:
                self.shadow[n] = (1.0 - self.decay) * p.data + self.decay * self.shadow[n]


class _Toy(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.body = torch.nn.Sequential(torch.nn.Conv2d(3, 16, 3), torch.nn.GELU(), torch.nn.Flatten(),
                                        torch.nn.Linear(16 * 6 * 6, 37), torch.nn.Linear(37, 5))
        self.unused = torch.nn.Linear(7, 3)      # never receives a gradient (the aux heads)

    def forward(self, x):
        return self.body(x)

    def __getitem__(self, i):
        return self.body[i]


def _toy_model(dev, seed):
    torch.manual_seed(seed)
    m = _Toy().to(dev)
    m[0].to(memory_format=torch.channels_last)              # a channels_last parameter
    return m


@pytest.mark.parametrize("use_scaler", [True, False])
def test_fused_optim_step_matches_reference_sequence(cnx_lib, use_scaler):
    """unscale_ + clip_grad_norm_ + AdamW.step + scaler.update + EMA.update (reference loop
    :226-237) against FusedOptimStep on identical gradients, including a step with an inf gradient
    (skipped update, halved scale) — 12 optimizer steps."""
    from image_classification_b200 import ema as ema_mod
    from image_classification_b200.optim_step import FusedOptimStep
    dev = _dev()
    ref, mine = _toy_model(dev, 3), _toy_model(dev, 3)
    opt_r = torch.optim.AdamW(ref.parameters(), lr=3e-3, weight_decay=1e-2)
    opt_m = torch.optim.AdamW(mine.parameters(), lr=3e-3, weight_decay=1e-2)
    sc_r = torch.amp.GradScaler("cuda", init_scale=1024.0, growth_interval=4) if use_scaler else None
    sc_m = torch.amp.GradScaler("cuda", init_scale=1024.0, growth_interval=4) if use_scaler else None
    ema_r, ema_m = _RefEMA(ref, 0.97), ema_mod.EMAB200(mine, 0.97)
    fused = FusedOptimStep(mine, opt_m, sc_m, ema_m, max_norm=0.05)      # small: clipping is active
    sched_r = torch.optim.lr_scheduler.LambdaLR(opt_r, lambda s: 1.0 / (1 + s))
    sched_m = torch.optim.lr_scheduler.LambdaLR(opt_m, lambda s: 1.0 / (1 + s))
    gen = torch.Generator(device="cpu").manual_seed(0)
    for step in range(12):
        x = torch.randn(4, 3, 8, 8, generator=gen).to(dev)
        y = torch.randn(4, 5, generator=gen).to(dev)
        for model, sc in ((ref, sc_r), (mine, sc_m)):
            loss = ((model(x) - y) ** 2).mean()
            (sc.scale(loss) if sc is not None else loss).backward()
        if step == 5:                       # overflow: both paths must skip and back off
            for model in (ref, mine):
                model[3].weight.grad[2, 3] = float("inf")
        # reference order
        if sc_r is not None:
            sc_r.unscale_(opt_r)
        norm_r = torch.nn.utils.clip_grad_norm_(ref.parameters(), 0.05)
        if sc_r is not None:
            sc_r.step(opt_r)
            sc_r.update()
        else:
            if step != 5:
                opt_r.step()                # (without a scaler the reference would take a NaN step)
        opt_r.zero_grad()
        sched_r.step()
        ema_r.update()
        # fused
        fused.step()
        opt_m.zero_grad()
        sched_m.step()
        if step != 5:
            assert torch.allclose(fused.total_norm, norm_r, rtol=1e-5), (step, fused.total_norm, norm_r)
        assert float(fused.found_inf) == (1.0 if step == 5 else 0.0)
        if use_scaler:
            assert sc_m.get_scale() == sc_r.get_scale(), step
        for (n, pr), (_, pm) in zip(ref.named_parameters(), mine.named_parameters()):
            assert torch.allclose(pm, pr, rtol=2e-5, atol=1e-7), (step, n, (pm - pr).abs().max())
            assert torch.allclose(ema_m.shadow[n], ema_r.shadow[n], rtol=2e-5, atol=1e-7), (step, n)
    # optimizer state stays torch-shaped: step count, moments
    for pr, pm in zip(ref.parameters(), mine.parameters()):
        if pr in opt_r.state:
            assert float(opt_m.state[pm]["step"]) == float(opt_r.state[pr]["step"]) == 11.0
            assert torch.allclose(opt_m.state[pm]["exp_avg"], opt_r.state[pr]["exp_avg"], rtol=1e-4, atol=1e-8)
            assert torch.allclose(opt_m.state[pm]["exp_avg_sq"], opt_r.state[pr]["exp_avg_sq"], rtol=1e-4, atol=1e-10)
    opt_m.state_dict()                      # still serialisable


def test_fused_optim_step_rejects_unsupported(cnx_lib):
    from image_classification_b200 import _lib
    from image_classification_b200.optim_step import FusedOptimStep
    m = _toy_model(_dev(), 0)
    with pytest.raises(_lib.CnxError):
        FusedOptimStep(m, torch.optim.SGD(m.parameters(), lr=0.1))
    with pytest.raises(_lib.CnxError):
        FusedOptimStep(m, torch.optim.AdamW(m.parameters(), amsgrad=True))


# ----------------------------------------------------------------- stem / downsample conv
def test_colsum_kernel_and_lone_patch_conv_is_plain_conv2d(cnx_lib):
    """cnx_colsum_bf16 (bias gradient of the patchify GEMMs) vs fp64; a PatchConv2d called on its
    own is nn.Conv2d (the second, cuDNN-backed implementation of round 1 is gone)."""
    from image_classification_b200 import ops
    from image_classification_b200.convnext import PatchConv2d
    dev = _dev()
    assert PatchConv2d.forward is torch.nn.Conv2d.forward
    for M, C in ((1000, 96), (4 * 14 * 14, 1024), (7, 8)):
        t = torch.randn(M, C, device=dev).bfloat16()
        cs = ops.colsum_bf16(t)
        assert float((cs.double() - t.double().sum(0)).abs().max()) < 1e-3


def test_trunk_modules_off_path_is_loud(cnx_lib, monkeypatch):
    """a CUDA activation that misses a library path of the ConvNeXt trunk raises
    under CNX_STRICT=1 and warns once (naming the failed condition) otherwise — never silent."""
    import warnings
    from image_classification_b200 import _lib, convnext
    from image_classification_b200.convnext import (DownsampleB200, GlobalAvgPoolB200, LayerNorm2d,
                                                    PatchConv2d, StemB200)
    dev = _dev()
    stem = StemB200(PatchConv2d(3, 64, kernel_size=4, stride=4), LayerNorm2d(64, eps=1e-6)).to(dev)
    down = DownsampleB200(LayerNorm2d(64, eps=1e-6), PatchConv2d(64, 128, kernel_size=2, stride=2)).to(dev)
    ln, pool = LayerNorm2d(64, eps=1e-6).to(dev), GlobalAvgPoolB200(1)
    img = torch.randn(2, 3, 32, 32, device=dev)
    f32 = torch.randn(2, 64, 8, 8, device=dev).contiguous(memory_format=torch.channels_last)
    bf = f32.bfloat16()
    cases = [
        ("StemB200", lambda: stem(img), "autocast"),                               # no autocast
        ("DownsampleB200", lambda: down(bf), "autocast"),                          # no autocast
        ("LayerNorm2d", lambda: ln(f32), "not bf16"),                              # fp32 stream
        ("GlobalAvgPoolB200", lambda: pool(f32), "not bf16"),
        ("LayerNorm2d", lambda: ln(bf.contiguous()), "channels_last"),             # NCHW-contiguous
    ]
    monkeypatch.setenv("CNX_STRICT", "1")
    for who, fn, frag in cases:
        with pytest.raises(_lib.CnxError) as ei:
            fn()
        assert who in str(ei.value) and frag in str(ei.value), str(ei.value)
    with torch.autocast("cuda", dtype=torch.float16):                             # fp16 autocast
        with pytest.raises(_lib.CnxError):
            stem(img)
        with pytest.raises(_lib.CnxError):
            down(bf)
    monkeypatch.setenv("CNX_STRICT", "0")
    convnext._warned.clear()
    for who, fn, frag in (cases[0], cases[2], cases[3]):   # (bf16 input + fp32 conv needs autocast)
        with warnings.catch_warnings(record=True) as rec:
            warnings.simplefilter("always")
            y = fn()
            fn()
        msgs = [str(w.message) for w in rec if who in str(w.message)]
        assert len(msgs) == 1 and frag in msgs[0], (who, msgs)     # once, with the reason
        assert torch.isfinite(y.float()).all()
    # the same modules on the path: no warning, library launches
    monkeypatch.setenv("CNX_STRICT", "1")
    n0 = _lib.launch_count
    with torch.autocast("cuda", dtype=torch.bfloat16):
        y = down(stem(img))
        pool(ln(bf))
    assert _lib.launch_count - n0 >= 6 and y.shape == (2, 128, 4, 4)


# ------------------------------------------------------------------------- input pipeline
def test_device_prefetcher_and_async_scalar_reader(cnx_lib):
    from image_classification_b200.prefetch import AsyncScalarReader, DevicePrefetcher
    dev = _dev()
    gen = torch.Generator().manual_seed(0)
    host = [(torch.randn(8, 3, 32, 32, generator=gen).pin_memory(),
             torch.randint(0, 44, (8,), generator=gen).pin_memory()) for _ in range(5)]
    reader = AsyncScalarReader(dev, depth=2)
    got, seen = [], 0
    for i, (x, y) in enumerate(DevicePrefetcher(iter(host), dev)):
        assert x.is_cuda and y.is_cuda
        # a long kernel chain on the compute stream while the next copy is in flight
        z = x
        for _ in range(20):
            z = z * 1.0001
        assert torch.equal(x.cpu(), host[i][0]) and torch.equal(y.cpu(), host[i][1])
        got += reader.push(z.sum() * 0 + float(i))
        seen += 1
    got += reader.drain()
    assert seen == 5 and got == [0.0, 1.0, 2.0, 3.0, 4.0]


@pytest.mark.parametrize("cfg", [(128, 256, 28, True), (256, 512, 14, False), (96, 192, 12, True)])
def test_downsample_b200_matches_torch_sequential(cnx_lib, cfg):
    """LayerNorm2d -> 2x2 s2 conv as cnx_layernorm_patch_fwd + tcgen05 GEMMs vs the same modules
    run by PyTorch in fp32/fp64: output, input gradient, all four parameter gradients."""
    from image_classification_b200.convnext import DownsampleB200, LayerNorm2d, PatchConv2d
    cin, cout, hw, cl = cfg
    dev = _dev()
    torch.manual_seed(cin)
    mine = DownsampleB200(LayerNorm2d(cin, eps=1e-6), PatchConv2d(cin, cout, kernel_size=2, stride=2)).to(dev)
    with torch.no_grad():
        mine[0].weight.copy_(1 + 0.2 * torch.randn(cin))
        mine[0].bias.copy_(0.1 * torch.randn(cin))
        mine[1].weight.copy_(0.05 * torch.randn_like(mine[1].weight))
        mine[1].bias.copy_(0.1 * torch.randn(cout))
    if cl:
        mine[1].to(memory_format=torch.channels_last)
    x = torch.randn(3, cin, hw, hw, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    g = torch.randn(3, cout, hw // 2, hw // 2, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    xm = x.clone().requires_grad_()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        ym = mine(xm)
    assert ym.dtype == torch.bfloat16 and ym.shape == g.shape
    ym.backward(g)
    # fp64 reference of the same two layers
    xr = x.double().requires_grad_()
    w = {k: v.detach().double() for k, v in mine.state_dict().items()}
    for v in w.values():
        v.requires_grad_()
    yn = torch.nn.functional.layer_norm(xr.permute(0, 2, 3, 1), (cin,), w["0.weight"], w["0.bias"], 1e-6)
    yr = torch.nn.functional.conv2d(yn.permute(0, 3, 1, 2), w["1.weight"], w["1.bias"], stride=2)
    yr.backward(g.double())
    rel = lambda a, b: float((a.double() - b).abs().max() / b.abs().max())
    assert rel(ym, yr.detach()) < 2e-2
    assert rel(xm.grad, xr.grad) < 2e-2
    assert mine[1].weight.grad.stride() == mine[1].weight.stride()
    for k, p in mine.named_parameters():
        assert rel(p.grad, w[k].grad) < 2e-2, (k, rel(p.grad, w[k].grad))


def test_stem_b200_matches_torch_sequential(cnx_lib):
    from image_classification_b200.convnext import LayerNorm2d, PatchConv2d, StemB200
    dev = _dev()
    torch.manual_seed(7)
    mine = StemB200(PatchConv2d(3, 128, kernel_size=4, stride=4), LayerNorm2d(128, eps=1e-6)).to(dev)
    mine[0].to(memory_format=torch.channels_last)
    with torch.no_grad():
        mine[0].weight.copy_(0.1 * torch.randn_like(mine[0].weight))
        mine[0].bias.copy_(0.1 * torch.randn(128))
        mine[1].weight.copy_(1 + 0.2 * torch.randn(128))
    x = torch.randn(4, 3, 64, 96, device=dev)
    g = torch.randn(4, 128, 16, 24, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        ym = mine(x)
    assert ym.dtype == torch.bfloat16 and ym.is_contiguous(memory_format=torch.channels_last)
    ym.backward(g)
    w = {k: v.detach().double().requires_grad_() for k, v in mine.state_dict().items()}
    xb = x.bfloat16().double()                       # the kernels see bf16 pixels
    yc = torch.nn.functional.conv2d(xb, w["0.weight"], w["0.bias"], stride=4)
    yr = torch.nn.functional.layer_norm(yc.permute(0, 2, 3, 1), (128,), w["1.weight"], w["1.bias"], 1e-6)
    yr = yr.permute(0, 3, 1, 2)
    yr.backward(g.double())
    rel = lambda a, b: float((a.double() - b).abs().max() / b.abs().max())
    assert rel(ym, yr.detach()) < 2e-2
    for k, p in mine.named_parameters():
        assert rel(p.grad, w[k].grad) < 2e-2, (k, rel(p.grad, w[k].grad))
    # the patchify permutation itself is exact
    from image_classification_b200 import ops
    P = ops.patchify_nchw_f32(x, 4)
    ref = x.bfloat16().view(4, 3, 16, 4, 24, 4).permute(0, 2, 4, 3, 5, 1).reshape(4 * 16 * 24, 48)
    assert torch.equal(P, ref)


# ------------------------------------------------------- timm batch-mode Mixup
@pytest.mark.parametrize("cfg", [dict(mixup_alpha=0.8, cutmix_alpha=0.0), dict(mixup_alpha=0.0, cutmix_alpha=1.0),
                                 dict(mixup_alpha=0.2, cutmix_alpha=1.0, switch_prob=0.5, prob=0.7)])
def test_timm_batch_mixup_bit_exact_vs_oracle(cnx_lib, cfg):
    from image_classification_b200 import mix
    dev = _dev()
    mixer = mix.TimmBatchMixup(label_smoothing=0.1, num_classes=44, **cfg)
    modes = set()
    for seed in range(8):
        gen = torch.Generator().manual_seed(seed)
        x = torch.randn(6, 3, 32, 48, generator=gen)
        y = torch.randint(0, 44, (6,), generator=gen)
        xr, tr, lam = mix_oracle.timm_batch_mixup_ref(x, y, label_smoothing=0.1, num_classes=44,
                                                       rng=np.random.RandomState(seed), **cfg)
        xm, tm = mixer(x.to(dev), y.to(dev), rng=np.random.RandomState(seed))
        assert torch.equal(xm.cpu(), xr), (seed, lam)
        assert torch.equal(tm.cpu(), tr), (seed, lam)
        assert abs(float(tm.sum(1).mean()) - 1.0) < 1e-5
        modes.add("pass" if lam == 1.0 else "mix")
    assert "mix" in modes
    with pytest.raises(Exception):
        mixer(torch.zeros(5, 3, 8, 8, device=dev), torch.zeros(5, dtype=torch.long, device=dev))


# ------------------------------------------------------------------ head: pool (+norm), aux heads
@pytest.mark.parametrize("shape", [(128, 7, 7, 1024), (16, 14, 14, 512), (5, 28, 28, 256),
                                   (3, 5, 3, 96), (2, 1, 1, 64), (4, 24, 24, 384)])
def test_avgpool_nhwc_fwd_bwd_vs_fp64(cnx_lib, shape):
    """cnx_avgpool_nhwc_fwd/_bwd against the definition in fp64 (bf16 rounding of the result only)."""
    from image_classification_b200 import ops
    dev = _dev()
    B, H, W, C = shape
    torch.manual_seed(B * 131 + C)
    x = (torch.randn(B, H, W, C, device=dev) + 0.5).bfloat16()
    y = ops.avgpool_nhwc_fwd(x)
    ref = x.double().mean(dim=(1, 2))
    assert y.shape == (B, C) and y.dtype == torch.bfloat16
    assert torch.equal(y, ref.to(torch.bfloat16)) or \
        float((y.double() - ref).abs().max()) <= 2 ** -8 * float(ref.abs().max())
    g = torch.randn(B, C, device=dev).bfloat16()
    dx = ops.avgpool_nhwc_bwd(g, H, W)
    want = (g.float() * (1.0 / (H * W))).to(torch.bfloat16)[:, None, None, :].expand(B, H, W, C)
    assert torch.equal(dx, want)


def test_avgpool_rejects_bad_shapes(cnx_lib):
    from image_classification_b200 import ops, _lib
    dev = _dev()
    with pytest.raises(_lib.CnxError):
        ops.avgpool_nhwc_fwd(torch.zeros(2, 3, 3, 12, device=dev, dtype=torch.bfloat16))   # C % 8
    with pytest.raises(_lib.CnxError):
        ops.avgpool_nhwc_fwd(torch.zeros(2, 3, 3, 16, device=dev))                         # fp32


def test_head_pool_norm_on_library_matches_torch(cnx_lib):
    """head.global_pool + head.norm on the channels_last bf16 stream (f-1) vs the PyTorch ops in fp32."""
    from image_classification_b200 import _lib
    from image_classification_b200.convnext import _Head
    dev = _dev()
    torch.manual_seed(5)
    head = _Head(512, 44).to(dev)
    with torch.no_grad():
        head.norm.weight.copy_(torch.rand(512, device=dev) + 0.5)
        head.norm.bias.copy_(torch.randn(512, device=dev) * 0.1)
    x0 = torch.randn(8, 512, 7, 7, device=dev)
    xa = x0.bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_()
    xb = xa.detach().float().requires_grad_()
    n0 = _lib.launch_count
    with torch.autocast("cuda", dtype=torch.bfloat16):
        ya = head(xa)
    assert _lib.launch_count - n0 == 2            # pool + layernorm rows
    p = xb.mean(dim=(2, 3))
    yb = torch.nn.functional.linear(
        torch.nn.functional.layer_norm(p, (512,), head.norm.weight, head.norm.bias, 1e-6),
        head.fc.weight, head.fc.bias)
    g = torch.randn_like(yb)
    ya.backward(g.to(ya.dtype))
    yb.backward(g)
    assert _rel(ya.float(), yb) < 2e-2
    assert _rel(xa.grad.float(), xb.grad) < 3e-2
    assert xa.grad.is_contiguous(memory_format=torch.channels_last)


def test_deep_supervision_wrapper_mirrors_reference(cnx_lib):
    """DeepSupervisionModelB200 = the reference's DeepSupervisionModel (:96-152): [main] + 3 aux
    outputs, aux heads get no gradient from the loop's outputs[0] loss; compute_aux=False drops them."""
    from image_classification_b200.convnext import ConvNeXtTimmLike, DeepSupervisionModelB200
    dev = _dev()
    torch.manual_seed(2)
    base = ConvNeXtTimmLike((1, 1, 1, 1), (64, 128, 256, 512), 44).to(dev)
    model = DeepSupervisionModelB200(base, 44).to(dev)
    assert [c[2].in_features for c in model.aux_classifiers] == [128, 256, 512]
    assert sorted(k.split(".")[0] for k in model.state_dict()).count("aux_classifiers") == 6
    x = torch.randn(4, 3, 64, 64, device=dev)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        outs = model(x)
        plain = base(x)
    assert isinstance(outs, list) and len(outs) == 4 and all(o.shape == (4, 44) for o in outs)
    assert torch.equal(outs[0], plain)
    # aux head i = Linear(mean over pixels of stage i+1's output)
    feats = []
    with torch.autocast("cuda", dtype=torch.bfloat16):
        h = base.stem(x)
        for i, st in enumerate(base.stages):
            h = st(h)
            if i > 0:
                feats.append(h)
    for f, clf, o in zip(feats, model.aux_classifiers, outs[1:]):
        want = torch.nn.functional.linear(f.float().mean(dim=(2, 3)), clf[2].weight, clf[2].bias)
        assert _rel(o.float(), want) < 2e-2
    outs[0].float().sum().backward()
    assert all(p.grad is None for p in model.aux_classifiers.parameters())
    assert all(p.grad is not None for p in base.parameters())
    model.compute_aux = False
    with torch.autocast("cuda", dtype=torch.bfloat16):
        only = model(x)
    assert len(only) == 1 and torch.equal(only[0], plain)
