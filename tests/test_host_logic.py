"""CPU-side checks: host logic, layering rules, and the N>1 gradient exchange on gloo (world 2)."""
import os
import re
import subprocess
import sys

import pytest
import numpy as np
import torch

from conftest import ROOT


def test_product_never_imports_the_oracle():
    """A product path that routes through oracle/ voids every parity claim."""
    pkg = os.path.join(ROOT, "image_classification_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, re.M), f
                assert "/root/reference" not in text or f.endswith(".py") or "//" in text


def test_nothing_on_the_gpu_path_reads_the_reference_tree():
    """/root/reference does not exist on the GPU box: only oracle/ref_loader.py and
    oracle/gen_golden.py (container-only tools) may open it; citations in comments are fine."""
    for rel in ("bench.py", "__graft_entry__.py", "image_classification_b200/mix.py",
                "image_classification_b200/ema.py", "image_classification_b200/train_step.py"):
        text = open(os.path.join(ROOT, rel)).read()
        assert "ref_loader" not in text and "load_reference" not in text, rel


def test_ops_fail_loudly_without_cuda():
    from image_classification_b200 import _lib, ops
    from image_classification_b200.block import ConvNeXtBlockB200
    from image_classification_b200.loss import SoftTargetCrossEntropyB200
    with pytest.raises(_lib.CnxError):
        ops.dwconv_ln_fwd(torch.zeros(1, 4, 4, 64, dtype=torch.bfloat16), torch.zeros(49, 64),
                          torch.zeros(64))
    with pytest.raises(_lib.CnxError):
        ConvNeXtBlockB200(64)(torch.zeros(1, 64, 4, 4))
    with pytest.raises(_lib.CnxError):
        SoftTargetCrossEntropyB200(0.1)(torch.zeros(2, 5), torch.zeros(2, 5))


def test_skeleton_matches_timm_attribute_contract():
    """What the reference touches: stem, stages[i].blocks[-1].conv_dw.out_channels
    (train_advanced_v4.py:106-111), head.{global_pool,norm,flatten,fc} (:140-143)."""
    from image_classification_b200.convnext import create_convnext
    from oracle.block_oracle import OracleConvNeXtBlock
    m = create_convnext("convnext_tiny", 44, block_cls=OracleConvNeXtBlock)
    assert [s.blocks[-1].conv_dw.out_channels for s in m.stages] == [96, 192, 384, 768]
    assert [len(s.blocks) for s in m.stages] == [3, 3, 9, 3]
    for attr in ("global_pool", "norm", "flatten", "fc"):
        assert hasattr(m.head, attr)
    x = torch.randn(2, 3, 64, 64)
    feats = m.stages(m.stem(x))
    out = m.head.fc(m.head.flatten(m.head.norm(m.head.global_pool(feats))))   # the :140-143 path
    assert torch.allclose(out, m(x)) and out.shape == (2, 44)
    n_params = sum(p.numel() for p in create_convnext("convnext_base", 44,
                                                       block_cls=OracleConvNeXtBlock).parameters())
    assert n_params == 87_611_564          # SURVEY.md §8 a1 (probe on the torchvision equivalent)


def test_skeleton_equals_torchvision_convnext():
    """The oracle model (skeleton + oracle block) reproduces torchvision's convnext_tiny."""
    import torchvision
    from image_classification_b200.convnext import create_convnext
    from oracle.block_oracle import OracleConvNeXtBlock, convert_torchvision_convnext
    torch.manual_seed(0)
    tv = torchvision.models.convnext_tiny(num_classes=44, stochastic_depth_prob=0.0).eval()
    m = create_convnext("convnext_tiny", 44, block_cls=OracleConvNeXtBlock).eval()
    missing = m.load_state_dict(convert_torchvision_convnext(tv.state_dict()), strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    x = torch.randn(2, 3, 64, 64)
    with torch.no_grad():
        assert torch.allclose(m(x), tv(x), atol=1e-5, rtol=1e-4)


DDP_WORKER = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.environ["CNX_ROOT"])
from image_classification_b200.ddp import BucketedGradReducer
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
torch.manual_seed(0)
model = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.GELU(), torch.nn.Linear(32, 8))
unused = torch.nn.Linear(4, 4)            # like the reference's discarded aux heads
params = list(model.parameters()) + list(unused.parameters())
red = BucketedGradReducer(params, bucket_bytes=1024)
g = torch.Generator().manual_seed(7)
X = torch.randn(2, 2, 6, 16, generator=g)      # [micro, rank, b, d]
Y = torch.randn(2, 2, 6, 8, generator=g)
def micro(i, r, m):
    return ((m(X[i, r]) - Y[i, r]) ** 2).mean()
# discover_unused after zero_grad (every .grad None) must refuse instead of dropping everything
try:
    red.discover_unused(); raise SystemExit("discover_unused accepted a model without gradients")
except RuntimeError:
    pass
assert len(red.params) == 6 and not red.discovered
# the first synchronised step finds the unused parameters by itself: the incomplete bucket is
# reduced with zeros standing in, then the buckets are rebuilt without the two unused tensors
red.set_sync(True); micro(0, rank, model).backward(); red.finish()
assert red.discovered and len(red.params) == 4 and len(red.dropped) == 2 and len(red.buckets) >= 2
assert red.reduced_steps == 1
m0 = [torch.zeros_like(p) for p in model.parameters()]
for r in range(world):
    m2 = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.GELU(), torch.nn.Linear(32, 8))
    m2.load_state_dict(model.state_dict())
    micro(0, r, m2).backward()
    for a, p in zip(m0, m2.parameters()): a += p.grad / world
for a, p in zip(m0, model.parameters()):
    assert torch.allclose(a, p.grad, atol=1e-6), ("discovery step", rank, (a - p.grad).abs().max())
for p in params: p.grad = None
# accumulate 2 micro-batches; exchange on the boundary only
red.set_sync(False); micro(0, rank, model).backward()
g_before = [p.grad.clone() for p in model.parameters()]
red.set_sync(True); micro(1, rank, model).backward(); red.finish()
# single-process truth: mean over ranks of (sum over micro-batches of local grads)
ref = [torch.zeros_like(p) for p in model.parameters()]
for r in range(world):
    m2 = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.GELU(), torch.nn.Linear(32, 8))
    m2.load_state_dict(model.state_dict())
    (micro(0, r, m2) + micro(1, r, m2)).backward()
    for a, p in zip(ref, m2.parameters()): a += p.grad / world
for a, p in zip(ref, model.parameters()):
    assert torch.allclose(a, p.grad, atol=1e-6), (rank, (a - p.grad).abs().max())
assert all(p.grad is None for p in unused.parameters())
# gradients that appear WITHOUT the hooks firing (what a CUDA-graph replay leaves in the static
# .grad buffers, graphed.GraphedForwardBackward): finish() sends every bucket itself
local = [torch.full_like(p, float(rank + 1)) for p in model.parameters()]
for p, gl in zip(model.parameters(), local): p.grad = gl.clone()
red.set_sync(True); red.finish()
for p in model.parameters():
    assert torch.allclose(p.grad, torch.full_like(p, 1.5)), rank      # mean of 1 and 2
# a known parameter without a gradient on a synchronised step: loud, not a silently skipped bucket
for p in model.parameters(): p.grad = None
red.set_sync(True)
try:
    list(model.parameters())[0].grad = None
    for p in list(model.parameters())[1:]: p.grad = torch.ones_like(p)
    red.finish(); raise SystemExit("finish() accepted an incomplete bucket after discovery")
except RuntimeError:
    pass
# a dropped parameter that does receive a gradient on a synchronised step: loud as well
red2 = BucketedGradReducer(params, bucket_bytes=1024)
for p in params: p.grad = None
red2.set_sync(True); micro(0, rank, model).backward(); red2.finish()
try:
    red2.set_sync(True); unused(torch.ones(1, 4)).sum().backward()
    raise SystemExit("a dropped parameter's gradient went unnoticed")
except RuntimeError:
    pass
dist.barrier(); dist.destroy_process_group()
print("rank", rank, "ok")
'''


def test_bucketed_grad_reducer_world2_gloo(tmp_path):
    script = tmp_path / "ddp_worker.py"
    script.write_text(DDP_WORKER)
    env = dict(os.environ, CNX_ROOT=ROOT, MASTER_ADDR="127.0.0.1", MASTER_PORT="29571",
               WORLD_SIZE="2", OMP_NUM_THREADS="1")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(2)]
    outs = [p.communicate(timeout=180)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, o
        assert f"rank {r} ok" in o


def test_bench_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference",
                          "--gpus", "2", "--steps", "1", "--warmup", "0"], env=env,
                         capture_output=True, text=True, timeout=120)
    assert res.returncode == 0 and res.stdout.strip() == ""


def test_timm_batch_mixup_planner_follows_oracle_rng_order():
    """Host planner of TimmBatchMixup vs the oracle restatement of timm's Mixup(mode='batch'):
    same RNG consumption, lambda, box and mode for every seed (no GPU needed)."""
    import numpy as np
    import torch
    from image_classification_b200 import mix
    from oracle import mix_oracle
    for cfg in (dict(mixup_alpha=0.8, cutmix_alpha=0.0), dict(mixup_alpha=0.0, cutmix_alpha=1.0),
                dict(mixup_alpha=0.2, cutmix_alpha=1.0, prob=0.7, switch_prob=0.3)):
        for seed in range(12):
            x = torch.zeros(4, 3, 16, 24)
            x[:, :, :, :] = torch.arange(4).view(4, 1, 1, 1).float()
            y = torch.arange(4)
            r1, r2 = np.random.RandomState(seed), np.random.RandomState(seed)
            xr, tr, lam = mix_oracle.timm_batch_mixup_ref(x, y, num_classes=44, rng=r1, **cfg)
            plan = mix.plan_timm_batch_mixup(4, 16, 24, rng=r2, **cfg)
            assert r1.rand() == r2.rand()                      # identical RNG consumption
            assert list(plan["partner"]) == [3, 2, 1, 0]
            if lam == 1.0:          # no mixing drawn, or a cut box clipped to zero area
                empty = (plan["bbx2"] - plan["bbx1"]) * (plan["bby2"] - plan["bby1"]) == 0
                assert ((plan["mode"] == mix.MODE_PASS) | ((plan["mode"] == mix.MODE_CUTMIX) & empty)).all()
                assert (plan["lam"] == 1.0).all()
                continue
            assert (plan["lam"] == np.float32(lam)).all()
            if plan["mode"][0] == mix.MODE_CUTMIX:
                pasted = (xr[0, 0] == 3.0)                     # sample 0 received pixels of sample 3
                box = np.zeros((16, 24), dtype=bool)
                box[plan["bby1"][0]:plan["bby2"][0], plan["bbx1"][0]:plan["bbx2"][0]] = True
                assert np.array_equal(pasted.numpy(), box)


def test_swap_trunk_keeps_parameters_names_and_results():
    """swap_trunk (f-1): stem / downsample / head pool + norm of a foreign timm-style model become
    the B200 modules around the SAME Parameter objects; on CPU they fall back to the PyTorch ops, so
    the logits must not change at all."""
    from image_classification_b200.convnext import (DownsampleB200, GlobalAvgPoolB200, LayerNorm2d,
                                                    StemB200, swap_trunk)
    from _foreign import _foreign_convnext
    torch.manual_seed(0)
    m = _foreign_convnext((1, 1, 1, 1), (16, 32, 64, 128)).eval()
    x = torch.randn(2, 3, 64, 64)
    before = m(x)
    keys = list(m.state_dict())
    params = {n: p for n, p in m.named_parameters()}
    assert swap_trunk(m) == 1 + 3 + 2
    assert isinstance(m.stem, StemB200) and isinstance(m.head.global_pool, GlobalAvgPoolB200)
    assert isinstance(m.head.norm, LayerNorm2d)
    assert all(isinstance(st.downsample, DownsampleB200) for st in m.stages[1:])
    assert list(m.state_dict()) == keys
    assert all(p is params[n] for n, p in m.named_parameters())
    assert torch.equal(m(x), before)
    assert swap_trunk(m) == 0                                     # idempotent


@pytest.mark.parametrize("k,stride,H,W", [(3, 1, 7, 9), (3, 2, 8, 8), (3, 2, 9, 7), (5, 1, 6, 11),
                                          (5, 2, 10, 13), (5, 2, 9, 9), (3, 1, 1, 1), (5, 2, 2, 3)])
def test_dwsep_oracle_pinned_to_torch_conv2d(k, stride, H, W):
    """oracle/dwsep_oracle.py against torch.nn.functional.conv2d (+ autograd) in float64: the op
    timm's efficientnet blocks call (previous/V3-efb2/train_advanced_v3.0.py:330)."""
    from oracle import dwsep_oracle as o
    torch.manual_seed(k * 100 + stride * 10 + H)
    B, C = 2, 6
    x = torch.randn(B, C, H, W, dtype=torch.float64, requires_grad=True)
    w = torch.randn(C, 1, k, k, dtype=torch.float64, requires_grad=True)
    y = torch.nn.functional.conv2d(x, w, stride=stride, padding=k // 2, groups=C)
    g = torch.randn_like(y)
    y.backward(g)
    xn = x.detach().permute(0, 2, 3, 1).numpy()
    wn = w.detach()[:, 0].permute(1, 2, 0).numpy()
    yo = o.dwconv_fwd_ref(xn, wn, stride)
    assert yo.shape == (B, o.out_size(H, k, stride), o.out_size(W, k, stride), C)
    np.testing.assert_allclose(yo, y.detach().permute(0, 2, 3, 1).numpy(), rtol=1e-12, atol=1e-12)
    dxo, dwo = o.dwconv_bwd_ref(g.permute(0, 2, 3, 1).numpy(), xn, wn, stride)
    np.testing.assert_allclose(dxo, x.grad.permute(0, 2, 3, 1).numpy(), rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(dwo, w.grad[:, 0].permute(1, 2, 0).numpy(), rtol=1e-12, atol=1e-12)


def test_swap_depthwise_separable_on_cpu_keeps_parameters_and_results():
    """f-4: the swapped convolutions hold the same Parameter objects and, off the fast path (CPU),
    are plain nn.Conv2d: identical logits, identical state_dict keys."""
    import torchvision
    from image_classification_b200.efficientnet import (DepthwiseConv2dB200, PointwiseConv2dB200,
                                                        swap_depthwise_separable)
    torch.manual_seed(0)
    m = torchvision.models.efficientnet_b2(weights=None, num_classes=44).eval()
    x = torch.randn(1, 3, 64, 64)
    before = m(x)
    keys = list(m.state_dict())
    params = {n: p for n, p in m.named_parameters()}
    n_dw, n_pw = swap_depthwise_separable(m)
    # 23 depthwise; 45 expand / project / head 1x1 convs + the 4 squeeze-excite ones whose channel
    # counts are multiples of 8 (those stay on nn.Conv2d.forward at run time: [B,C,1,1] maps)
    assert (n_dw, n_pw) == (23, 49)
    assert sum(isinstance(c, DepthwiseConv2dB200) for c in m.modules()) == 23
    assert sum(isinstance(c, PointwiseConv2dB200) for c in m.modules()) == 49
    assert list(m.state_dict()) == keys
    assert all(p is params[n] for n, p in m.named_parameters())
    assert torch.equal(m(x), before)
    assert swap_depthwise_separable(m) == (0, 0)


@pytest.mark.parametrize("silu", [True, False])
def test_bn_act_oracle_pinned_to_torch_batch_norm(silu):
    """oracle bn_act_* against F.batch_norm(training=True) [-> F.silu] and autograd in float64,
    including nn.BatchNorm2d's running-buffer update."""
    from oracle import dwsep_oracle as o
    torch.manual_seed(3 + silu)
    M, C, eps, mom = 37, 5, 1e-5, 0.1
    x = (torch.randn(M, C, dtype=torch.float64) * 2 + 0.7).requires_grad_()
    w = torch.randn(C, dtype=torch.float64).requires_grad_()
    b = torch.randn(C, dtype=torch.float64).requires_grad_()
    rm, rv = torch.randn(C, dtype=torch.float64), torch.rand(C, dtype=torch.float64) + 0.5
    rm0, rv0 = rm.clone(), rv.clone()
    y = torch.nn.functional.batch_norm(x, rm, rv, w, b, True, mom, eps)
    if silu:
        y = torch.nn.functional.silu(y)
    g = torch.randn_like(y)
    y.backward(g)
    yo, mean, var = o.bn_act_fwd_ref(x.detach().numpy(), w.detach().numpy(), b.detach().numpy(), eps, silu)
    np.testing.assert_allclose(yo, y.detach().numpy(), rtol=1e-11, atol=1e-12)
    rmo, rvo = o.bn_running_update_ref(rm0.numpy(), rv0.numpy(), mean, var, M, mom)
    np.testing.assert_allclose(rmo, rm.numpy(), rtol=1e-12)
    np.testing.assert_allclose(rvo, rv.numpy(), rtol=1e-12)
    dx, dw, db = o.bn_act_bwd_ref(g.numpy(), x.detach().numpy(), w.detach().numpy(),
                                  b.detach().numpy(), eps, silu)
    np.testing.assert_allclose(dx, x.grad.numpy(), rtol=1e-9, atol=1e-11)
    np.testing.assert_allclose(dw, w.grad.numpy(), rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(db, b.grad.numpy(), rtol=1e-10, atol=1e-12)


def test_se_oracle_pinned_to_torchvision_squeeze_excitation():
    """oracle se_* against torchvision.ops.SqueezeExcitation(SiLU, Sigmoid) — the block of
    efficientnet_b2 — and autograd in float64."""
    import torchvision
    from oracle import dwsep_oracle as o
    torch.manual_seed(5)
    B, C, Cs, H, W = 3, 8, 2, 4, 5
    se = torchvision.ops.SqueezeExcitation(C, Cs, activation=torch.nn.SiLU).double()
    x = torch.randn(B, C, H, W, dtype=torch.float64, requires_grad=True)
    y = se(x)
    g = torch.randn_like(y)
    y.backward(g)
    w1, b1 = se.fc1.weight.detach().reshape(Cs, C).numpy(), se.fc1.bias.detach().numpy()
    w2, b2 = se.fc2.weight.detach().reshape(C, Cs).numpy(), se.fc2.bias.detach().numpy()
    xn = x.detach().permute(0, 2, 3, 1).reshape(B, H * W, C).numpy()
    gn = g.permute(0, 2, 3, 1).reshape(B, H * W, C).numpy()
    yo, _ = o.se_fwd_ref(xn, w1, b1, w2, b2)
    np.testing.assert_allclose(yo, y.detach().permute(0, 2, 3, 1).reshape(B, H * W, C).numpy(), rtol=1e-12)
    dx, dw1, db1, dw2, db2 = o.se_bwd_ref(gn, xn, w1, b1, w2, b2)
    np.testing.assert_allclose(dx, x.grad.permute(0, 2, 3, 1).reshape(B, H * W, C).numpy(), rtol=1e-10, atol=1e-13)
    np.testing.assert_allclose(dw1, se.fc1.weight.grad.reshape(Cs, C).numpy(), rtol=1e-10, atol=1e-13)
    np.testing.assert_allclose(db1, se.fc1.bias.grad.numpy(), rtol=1e-10, atol=1e-13)
    np.testing.assert_allclose(dw2, se.fc2.weight.grad.reshape(C, Cs).numpy(), rtol=1e-10, atol=1e-13)
    np.testing.assert_allclose(db2, se.fc2.bias.grad.numpy(), rtol=1e-10, atol=1e-13)


def test_swap_efficientnet_on_cpu_keeps_parameters_buffers_and_results():
    """f-4: conv + BatchNorm(+SiLU) + squeeze-excite modules swapped around the same Parameter and
    buffer objects; off the fast path (CPU) the model computes what it did before, in eval and in
    train mode (running statistics included)."""
    import copy
    import torchvision
    from image_classification_b200.efficientnet import BatchNormActB200, swap_efficientnet
    torch.manual_seed(0)
    m = torchvision.models.efficientnet_b2(weights=None, num_classes=44)
    for mod in m.modules():
        if isinstance(mod, (torchvision.ops.StochasticDepth, torch.nn.Dropout)):
            mod.p = 0.0
    ref = copy.deepcopy(m)
    keys = list(m.state_dict())
    params = {n: p for n, p in m.named_parameters()}
    bufs = {n: b for n, b in m.named_buffers()}
    counts = swap_efficientnet(m)
    assert counts == {"depthwise": 23, "pointwise": 49, "batchnorm": 69, "squeeze_excite": 23}
    assert sum(isinstance(c, BatchNormActB200) and c.emit_pool for c in m.modules()) == 23
    assert not any(type(c) is torch.nn.BatchNorm2d for c in m.modules())
    assert list(m.state_dict()) == keys
    assert all(p is params[n] for n, p in m.named_parameters())
    assert all(b is bufs[n] for n, b in m.named_buffers())
    x = torch.randn(2, 3, 64, 64)
    m.train(), ref.train()
    assert torch.allclose(m(x), ref(x), rtol=1e-5, atol=1e-6)
    for (n, a), (_, b) in zip(m.named_buffers(), ref.named_buffers()):
        assert torch.allclose(a.float(), b.float(), rtol=1e-5, atol=1e-7), n
    m.eval(), ref.eval()
    assert torch.allclose(m(x), ref(x), rtol=1e-5, atol=1e-6)
    again = swap_efficientnet(m)
    assert again["batchnorm"] == 0 and again["squeeze_excite"] == 0


def test_swap_efficientnet_timm_layout_on_cpu():
    """timm's layout (BatchNormAct2d modules outside nn.Sequential, conv_reduce / conv_expand
    squeeze-excite, DepthwiseSeparableConv / InvertedResidual blocks; stand-ins in tests/_foreign.py
    since timm is not installable): everything is swapped, state_dict keys / Parameter / buffer
    objects are kept, and off the fast path the model computes what it did before."""
    import copy
    from _foreign import TimmLikeEfficientNet
    from image_classification_b200.efficientnet import (BatchNormActB200, SqueezeExciteB200,
                                                        swap_efficientnet)
    torch.manual_seed(0)
    m = TimmLikeEfficientNet()
    ref = copy.deepcopy(m)
    keys = list(m.state_dict())
    params = {n: p for n, p in m.named_parameters()}
    bufs = {n: b for n, b in m.named_buffers()}
    counts = swap_efficientnet(m)
    assert counts == {"depthwise": 6, "pointwise": 13, "batchnorm": 18, "squeeze_excite": 6}
    assert not any(isinstance(c, torch.nn.BatchNorm2d) and not isinstance(c, BatchNormActB200)
                   for c in m.modules())
    assert sum(isinstance(c, SqueezeExciteB200) for c in m.modules()) == 6
    # the BatchNorm feeding each squeeze-excite block (bn1 of DepthwiseSeparableConv, bn2 of
    # InvertedResidual) hands over its pooled sums
    assert m.blocks[0][0].bn1.emit_pool and m.blocks[1][0].bn2.emit_pool
    assert not m.blocks[1][0].bn1.emit_pool and not m.blocks[1][0].bn3.emit_pool
    assert m.blocks[1][0].bn3.act is None and m.blocks[1][0].bn2.act == "silu"
    assert list(m.state_dict()) == keys
    assert all(p is params[n] for n, p in m.named_parameters())
    assert all(b is bufs[n] for n, b in m.named_buffers())
    x = torch.randn(2, 3, 64, 64)
    m.train(), ref.train()
    assert torch.allclose(m(x), ref(x), rtol=1e-5, atol=1e-6)
    for (n, a), (_, b) in zip(m.named_buffers(), ref.named_buffers()):
        assert torch.allclose(a.float(), b.float(), rtol=1e-5, atol=1e-7), n
    m.eval(), ref.eval()
    assert torch.allclose(m(x), ref(x), rtol=1e-5, atol=1e-6)


def test_averaged_model_b200_interface_on_cpu():
    """AveragedModelB200 mirrors torch.optim.swa_utils.AveragedModel (the reference's SWA object,
    previous/V3-efb2/train_advanced_v3.0.py:445): same state_dict keys and forward; the update has
    no CPU path and says so."""
    import copy
    from torch.optim.swa_utils import AveragedModel
    from image_classification_b200 import _lib
    from image_classification_b200.ema import AveragedModelB200
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.SiLU(), torch.nn.Linear(16, 4))
    ours, ref = AveragedModelB200(model), AveragedModel(copy.deepcopy(model))
    assert list(ours.state_dict()) == list(ref.state_dict())
    assert int(ours.n_averaged) == 0
    x = torch.randn(3, 8)
    assert torch.equal(ours(x), ref(x))
    assert all(a is not b for a, b in zip(ours.module.parameters(), model.parameters()))
    with pytest.raises(_lib.CnxError):
        ours.update_parameters(model)
