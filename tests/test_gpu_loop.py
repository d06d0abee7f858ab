"""Row a6 on the device: the product's `TrainStep` (fused CUDA optimizer step on AND off, library
soft-target CE, library EMA) against tests/golden/loop_reference.npz — the output of the
reference's OWN `train_one_epoch` (/root/reference/train_advanced_v4.py:193-281), `EMA` (:62-94)
and schedule (:183-191), produced by oracle/gen_golden.py::gen_loop.  fp32 on the device, 1e-5.
Plus what the fused step must not break: the blocks' folded-weight cache and the
accumulation window's odd tail."""
import os
import warnings

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import block_oracle  # noqa: E402  (checker only)
from oracle.gen_golden import loop_fixture_inputs  # noqa: E402
from oracle.loop_oracle import TinyListModel  # noqa: E402

TOL = 1e-5


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a B200"
    return torch.device("cuda:0")


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("case,amp", [("plain", False), ("amp", True), ("plain_odd", False)])
def test_train_step_vs_reference_loop_fixture(cnx_lib, golden_dir, case, amp, fused):
    from image_classification_b200 import ema as ema_mod, loss as loss_mod, train_step
    dev = _dev()
    g = np.load(os.path.join(golden_dir, "loop_reference.npz"))
    model = TinyListModel()
    model.load_state_dict({k[len("init_"):]: torch.from_numpy(g[k]) for k in g.files
                           if k.startswith("init_")})
    model = model.to(dev)
    n = int(g[f"{case}_n_micro"])
    opt = torch.optim.AdamW(model.parameters(), lr=5e-3, weight_decay=1e-2)
    sched = train_step.get_cosine_schedule_with_warmup(opt, 1, 3, 1e-6)
    ema = ema_mod.EMAB200(model, 0.9)
    # AMP branch with a LIVE GradScaler (scale 2^16: scaling and unscaling are exact) and autocast
    # pointed at float32 so the arithmetic stays comparable at 1e-5
    scaler = torch.amp.GradScaler("cuda") if amp else None
    cfg = train_step.StepConfig(2, 1.0, 10, amp_dtype=torch.float32)
    step = train_step.TrainStep(model, loss_mod.SoftTargetCrossEntropyB200(0.1), opt, sched, scaler,
                                ema, cfg, use_fused_optim=fused)
    assert (step.fused is not None) == fused
    s = cnt = 0.0
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")            # "autocast target dtype float32: disabling"
        for i, (x, t) in enumerate(loop_fixture_inputs(n)):
            loss = step.micro_step(x.to(dev), t.to(dev), i)
            s += float(loss) * x.size(0)
            cnt += x.size(0)
            for k, v in model.named_parameters():
                ref = g[f"{case}_m{i}_{k}"]
                got = v.detach().cpu().numpy()
                assert np.allclose(got, ref, rtol=TOL, atol=TOL), (case, fused, i, k,
                                                                   np.abs(got - ref).max())
            assert abs(opt.param_groups[0]["lr"] - float(g[f"{case}_lrs"][i])) < 1e-15
    step.close()
    assert abs(s / cnt - float(g[f"{case}_mean_loss"])) < 1e-4
    for k, sh in ema.shadow.items():
        ref = g[f"{case}_shadow_{k}"]
        assert np.allclose(sh.cpu().numpy(), ref, rtol=TOL, atol=TOL), (case, fused, "shadow", k)
    for k, p in model.named_parameters():
        key = f"{case}_grad_{k}"
        if key in g.files:      # odd tail: the gradients stay in .grad (:225-231 did not run)
            assert p.grad is not None
            assert np.allclose(p.grad.cpu().numpy(), g[key], rtol=TOL, atol=TOL), (case, "grad", k)


def _block_pair(C, dev):
    from image_classification_b200.block import ConvNeXtBlockB200
    torch.manual_seed(0)
    ref = block_oracle.OracleConvNeXtBlock(C).to(dev)
    with torch.no_grad():
        for n, p in ref.named_parameters():
            if n == "gamma":
                p.copy_(torch.rand_like(p) * 0.9 + 0.1)
            elif p.dim() == 1 and n != "norm.weight":
                p.copy_(torch.randn_like(p) * 0.1)
            elif p.dim() > 1:
                p.mul_(3.0)
    prod = ConvNeXtBlockB200(C).to(dev).load_from(ref)
    return ref, prod


def _rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def test_block_forward_after_fused_step_uses_the_new_weights(cnx_lib):
    """FusedOptimStep writes parameters through raw pointers (torch's version counters do not
    move): the next forward must run on weights folded from the UPDATED parameters."""
    from image_classification_b200.optim_step import FusedOptimStep
    dev = _dev()
    ref, prod = _block_pair(128, dev)
    opt = torch.optim.AdamW(prod.parameters(), lr=5e-2, weight_decay=0.0)     # a step that matters
    fused = FusedOptimStep(prod, opt, None, None, max_norm=0.0)
    x = torch.randn(2, 128, 12, 12, device=dev).to(torch.bfloat16).contiguous(
        memory_format=torch.channels_last)
    y0 = prod(x)
    y0.float().square().mean().backward()
    before = {n: p.detach().clone() for n, p in prod.named_parameters()}
    fused.step()
    opt.zero_grad()
    moved = max(float((p.detach() - before[n]).abs().max()) for n, p in prod.named_parameters())
    assert moved > 1e-2
    y1 = prod(x)
    ref.load_state_dict(prod.state_dict())
    y_ref = ref(x.float())
    stale = block_oracle.OracleConvNeXtBlock(128).to(dev)
    stale.load_state_dict(before)
    y_stale = stale(x.float())
    assert _rel(y1.float() - x.float(), y_ref - x.float()) < 2e-2
    assert _rel(y1.float() - x.float(), y_stale - x.float()) > 0.1    # and NOT the old weights


def test_grad_window_odd_tail_flushes_into_grad(cnx_lib):
    """A window that never reaches its boundary (odd number of micro-batches) hands its
    accumulated block gradients to `.grad` at `TrainStep.close()`, where the reference's plain
    accumulation leaves them."""
    from image_classification_b200.block import GradWindow
    dev = _dev()
    ref, prod = _block_pair(64, dev)
    x = torch.randn(2, 64, 9, 7, device=dev)
    xb = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    ref(x).square().mean().backward()
    win = GradWindow(prod)
    win.begin()
    win.set_boundary(False)
    prod(xb).float().square().mean().backward()
    assert all(p.grad is None for p in prod.parameters())
    win.flush()
    win.detach()
    for (n, p), (_, q) in zip(prod.named_parameters(), ref.named_parameters()):
        assert p.grad is not None and _rel(p.grad, q.grad) < 3e-2, (n, _rel(p.grad, q.grad))
    # after the flush the window is closed: a plain backward works and accumulates on top
    prod(xb).float().square().mean().backward()
    for (n, p), (_, q) in zip(prod.named_parameters(), ref.named_parameters()):
        assert _rel(p.grad, 2 * q.grad) < 3e-2, n


def test_mbarrier_timeout_is_sticky_until_cleared(cnx_lib):
    """A recorded mbarrier timeout (results of that launch are garbage) makes every later entry
    point fail with CNX_ERR_LAUNCH until cnx_debug_timeout_flag() reads and clears it."""
    from image_classification_b200 import _lib, ops
    dev = _dev()
    x = torch.randn(64, 128, device=dev).bfloat16()
    ops.colsum_bf16(x)
    lib = _lib.load()
    # the flag page is created by the first TMA kernel launch
    ops.gemm_bf16(x, x)
    torch.cuda.synchronize()
    assert lib.cnx_debug_timeout_flag() == 0
    lib.cnx_debug_set_timeout_flag(1)
    try:
        with pytest.raises(_lib.CnxError) as ei:
            ops.colsum_bf16(x)
        assert "timed out" in str(ei.value)
        with pytest.raises(_lib.CnxError):
            ops.gemm_bf16(x, x)
    finally:
        assert lib.cnx_debug_timeout_flag() == 1      # reads and clears
    assert lib.cnx_debug_timeout_flag() == 0
    ops.colsum_bf16(x)
    torch.cuda.synchronize()


def test_graphed_train_step_matches_eager(cnx_lib):
    """graphed.GraphedTrainStep (two CUDA graphs per accumulate-2 window, weight fold recorded in
    the first, fold adjoint in the second) against the eager TrainStep on identical models and
    data: gradients of the window, then losses / parameters / EMA over further optimizer steps —
    which also shows that a replay folds the UPDATED weights."""
    import copy
    from image_classification_b200 import ema as ema_mod, loss as loss_mod, train_step
    from image_classification_b200.convnext import ConvNeXtTimmLike, DeepSupervisionModelB200
    from image_classification_b200.graphed import GraphedTrainStep
    dev = _dev()
    torch.manual_seed(5)
    base = DeepSupervisionModelB200(ConvNeXtTimmLike((1, 1, 2, 1), (64, 128, 256, 512), 44), 44).to(dev)
    with torch.no_grad():
        for n, p in base.named_parameters():
            if n.endswith("gamma"):
                p.copy_(torch.rand_like(p) * 0.9 + 0.1)
    models = [base, copy.deepcopy(base)]
    g = torch.Generator().manual_seed(9)
    xs = [torch.randn(4, 3, 64, 64, generator=g).to(dev) for _ in range(6)]
    ts = [torch.softmax(torch.randn(4, 44, generator=g) * 2, 1).to(dev) for _ in range(6)]
    steps = []
    for m in models:
        m.train()
        opt = torch.optim.AdamW(m.parameters(), lr=1e-3, weight_decay=1e-2)
        steps.append(train_step.TrainStep(m, loss_mod.SoftTargetCrossEntropyB200(0.1), opt, None,
                                          torch.amp.GradScaler("cuda"), ema_mod.EMAB200(m, 0.99),
                                          train_step.StepConfig(2, 1.0, 44)))
    eager, gts = steps[0], GraphedTrainStep(steps[1], (4, 3, 64, 64), (4, 44), dev)
    assert len(gts.graphs) == 2 and gts.launches[0] > gts.launches[1] - 20 and gts.launches[0] > 30
    # 1) gradients of one window, before any optimizer step
    for k in range(2):
        eager.forward_backward(xs[k], ts[k], k)
        gts.images.copy_(xs[k]); gts.targets.copy_(ts[k])
        gts.graphs[k].replay()
    for (n, p), gg in zip(models[0].named_parameters(), gts.grads):
        if p.grad is None:
            assert gg is None, n
            continue
        s = float(p.grad.abs().max().clamp_min(1e-30))
        assert float((p.grad - gg).abs().max()) / s < 1e-3, n        # atomics order only
    eager.optimizer_section()
    for p, gg in zip(gts.params, gts.grads):
        p.grad = gg
    gts.step.optimizer_section(reduce_grads=False)
    # 2) two more optimizer steps through the public calls
    for i in range(2, 6):
        le = eager.micro_step(xs[i], ts[i], i)
        lg = gts.micro_step(xs[i], ts[i], i)
        assert abs(float(le) - float(lg)) < 2e-3 * abs(float(le)), (i, float(le), float(lg))
    for (n, p), (_, q) in zip(models[0].named_parameters(), models[1].named_parameters()):
        assert float((p - q).abs().max()) < 5e-3 * float(p.abs().max().clamp_min(1e-3)) + 2e-3, n
    for k in steps[0].ema.shadow:
        assert torch.allclose(steps[0].ema.shadow[k], steps[1].ema.shadow[k], rtol=1e-2, atol=2e-3), k
    gts.close()
