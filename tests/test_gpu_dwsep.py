"""GPU parity for the depthwise-separable path (BASELINE config 5): the k x k
depthwise kernels against oracle/dwsep_oracle.py (pinned to torch's conv2d on the CPU side) at
small ragged shapes, against float64 conv2d on the device at EfficientNet-B2's own shapes, and
torchvision's efficientnet_b2 with its convolutions swapped against the untouched fp32 model.
bf16 tolerance 2e-2 of the tensor's max (the north-star's bf16 bound), written at each assert."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from oracle import dwsep_oracle  # noqa: E402  (checker only)

BF16_REL = 2e-2


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a B200"
    return torch.device("cuda:0")


def _rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _tap_major(w4):          # [C,1,k,k] -> [k*k, C] fp32
    C, _, k, _ = w4.shape
    return w4.reshape(C, k * k).t().contiguous().float()


@pytest.mark.parametrize("k,stride,B,H,W,C", [
    (3, 1, 2, 7, 9, 16), (3, 2, 2, 8, 8, 24), (3, 2, 1, 9, 7, 8), (5, 1, 2, 6, 11, 40),
    (5, 2, 2, 10, 13, 16), (5, 2, 1, 9, 9, 72), (3, 1, 1, 1, 1, 8), (5, 2, 3, 2, 3, 8),
    (3, 1, 1, 5, 4, 264), (5, 1, 1, 3, 17, 96)])
def test_dwconv_kxk_vs_oracle_small(cnx_lib, k, stride, B, H, W, C):
    """ragged strips (W not a multiple of 4), single pixels, odd sizes with stride 2, C not a
    power of two: bf16 inputs are exact in the oracle, so only the output rounding differs."""
    from image_classification_b200 import ops
    dev = _dev()
    rng = np.random.RandomState(k * 1000 + stride * 100 + H * 10 + C)
    x = torch.from_numpy(rng.randn(B, H, W, C).astype(np.float32)).bfloat16()
    w4 = torch.from_numpy(rng.randn(C, 1, k, k).astype(np.float32) * 0.3).bfloat16().float()
    wn = w4[:, 0].permute(1, 2, 0).double().numpy()                 # [k,k,C]
    y_ref = dwsep_oracle.dwconv_fwd_ref(x.double().numpy(), wn, stride)
    wt = _tap_major(w4).to(dev)
    y = ops.dwconv_kxk_fwd(x.to(dev), wt, k, stride)
    assert tuple(y.shape) == y_ref.shape
    assert _rel(y.cpu(), torch.from_numpy(y_ref)) < 2.0 ** -8            # one bf16 rounding
    dy = torch.from_numpy(rng.randn(*y_ref.shape).astype(np.float32)).bfloat16()
    dx_ref, dw_ref = dwsep_oracle.dwconv_bwd_ref(dy.double().numpy(), x.double().numpy(), wn, stride)
    dx, dwt = ops.dwconv_kxk_bwd(dy.to(dev), x.to(dev), wt, k, stride)
    assert _rel(dx.cpu(), torch.from_numpy(dx_ref)) < 2.0 ** -8
    assert _rel(dwt.cpu(), torch.from_numpy(dw_ref).reshape(k * k, C)) < 1e-5   # fp32 sums
    # accumulate mode adds on top; dx-only / dw-only calls
    dx2, dwt2 = ops.dwconv_kxk_bwd(dy.to(dev), x.to(dev), wt, k, stride, dwt_acc=dwt.clone())
    assert _rel(dwt2, 2 * dwt) < 1e-5 and torch.equal(dx2, dx)
    only_dx, none = ops.dwconv_kxk_bwd(dy.to(dev), x.to(dev), wt, k, stride, need_dw=False)
    assert none is None and torch.equal(only_dx, dx)
    none, only_dw = ops.dwconv_kxk_bwd(dy.to(dev), x.to(dev), wt, k, stride, need_dx=False)
    assert none is None and _rel(only_dw, dwt) < 1e-5


# (k, stride, C, H) of every depthwise convolution in efficientnet_b2 at 260x260
EFFNET_B2_DW = [(3, 1, 32, 130), (3, 1, 16, 130), (3, 2, 96, 130), (3, 1, 144, 65), (5, 2, 144, 65),
                (5, 1, 288, 33), (3, 2, 288, 33), (3, 1, 528, 17), (5, 1, 528, 17), (5, 1, 720, 17),
                (5, 2, 720, 17), (5, 1, 1248, 9), (3, 1, 1248, 9), (3, 1, 2112, 9)]


@pytest.mark.parametrize("k,stride,C,H", EFFNET_B2_DW)
def test_dwconv_kxk_effnet_b2_shapes_vs_conv2d_fp64(cnx_lib, k, stride, C, H):
    """Batch 48 (previous/V3-efb2/config.py:14) at the network's own shapes, against
    F.conv2d(groups=C) + autograd in float64 on the device."""
    from image_classification_b200 import ops
    dev = _dev()
    B = 48 if C * H * H <= 96 * 130 * 130 else 16
    torch.manual_seed(C + H)
    x = torch.randn(B, C, H, H, device=dev).bfloat16()
    w4 = (torch.randn(C, 1, k, k, device=dev) * 0.2).bfloat16().float()
    xr = x.double().requires_grad_()
    wr = w4.double().requires_grad_()
    yr = F.conv2d(xr, wr, stride=stride, padding=k // 2, groups=C)
    g = torch.randn_like(yr).bfloat16()
    yr.backward(g.double())
    xn = x.permute(0, 2, 3, 1).contiguous()
    wt = _tap_major(w4)
    y = ops.dwconv_kxk_fwd(xn, wt, k, stride)
    assert _rel(y.permute(0, 3, 1, 2), yr.detach()) < 2.0 ** -8
    dx, dwt = ops.dwconv_kxk_bwd(g.permute(0, 2, 3, 1).contiguous(), xn, wt, k, stride)
    assert _rel(dx.permute(0, 3, 1, 2), xr.grad) < 2.0 ** -8
    assert _rel(dwt.t().reshape(C, 1, k, k), wr.grad) < 1e-4          # fp32 atomics over 10^5 terms


def test_dwconv_kxk_rejects_bad_arguments(cnx_lib):
    from image_classification_b200 import ops, _lib
    dev = _dev()
    x = torch.zeros(1, 4, 4, 16, device=dev, dtype=torch.bfloat16)
    with pytest.raises(_lib.CnxError):
        ops.dwconv_kxk_fwd(x, torch.zeros(49, 16, device=dev), 7, 1)           # k = 7: dwconv7 path
    with pytest.raises(_lib.CnxError):
        ops.dwconv_kxk_fwd(x, torch.zeros(9, 16, device=dev), 3, 3)            # stride 3
    with pytest.raises(_lib.CnxError):
        ops.dwconv_kxk_fwd(x[..., :12].contiguous(), torch.zeros(9, 12, device=dev), 3, 1)   # C % 8
    with pytest.raises(_lib.CnxError):
        ops.dwconv_kxk_fwd(x.float(), torch.zeros(9, 16, device=dev), 3, 1)    # fp32 activations


@pytest.mark.parametrize("cfg", [(96, 24, 65, False), (144, 48, 33, False), (32, 16, 130, False),
                                 (528, 88, 17, True), (352, 1408, 9, False)])
def test_pointwise_conv_module_matches_conv2d(cnx_lib, cfg):
    """1x1 convolutions of the MBConv blocks (expand / project / head widths, N and K far from
    tile multiples) through the tcgen05 GEMM vs nn.Conv2d in fp32."""
    from image_classification_b200 import _lib
    from image_classification_b200.efficientnet import PointwiseConv2dB200
    dev = _dev()
    Cin, Cout, H, bias = cfg
    torch.manual_seed(Cin + Cout)
    conv = PointwiseConv2dB200(Cin, Cout, 1, bias=bias).to(dev)
    x0 = torch.randn(16, Cin, H, H, device=dev)          # >= PointwiseConv2dB200.min_rows pixels
    xa = x0.bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_()
    xb = xa.detach().float().requires_grad_()
    n0 = _lib.launch_count
    with torch.autocast("cuda", dtype=torch.bfloat16):
        ya = conv(xa)
    assert _lib.launch_count > n0 and ya.dtype == torch.bfloat16
    wq = conv.weight.detach().bfloat16().float()          # the GEMM sees bf16 weights (autocast)
    yb = F.conv2d(xb, wq, conv.bias)
    g = torch.randn_like(yb).bfloat16()
    ya.backward(g)
    gw, gb = conv.weight.grad.clone(), None if conv.bias is None else conv.bias.grad.clone()
    conv.weight.grad = None
    yb.backward(g.float())
    assert _rel(ya.float(), yb) < BF16_REL
    assert _rel(xa.grad.float(), xb.grad) < BF16_REL
    wref = torch.autograd.grad(F.conv2d(xb.detach(), conv.weight, conv.bias), conv.weight, g.float())[0]
    assert _rel(gw, wref) < 1e-3
    if bias:
        assert _rel(gb, g.float().sum(dim=(0, 2, 3))) < 1e-4


def test_efficientnet_b2_swapped_vs_fp32_model(cnx_lib):
    """torchvision efficientnet_b2 (same architecture as timm's, random init, 44 classes) with its
    23 depthwise and 45 large 1x1 convolutions on libcnx_b200 under bf16 autocast, against the
    untouched model in fp32: logits, loss and gradient norm within the bf16 bound."""
    import copy
    import torchvision
    from image_classification_b200 import _lib
    from image_classification_b200.efficientnet import swap_depthwise_separable
    dev = _dev()
    torch.manual_seed(0)
    ref = torchvision.models.efficientnet_b2(weights=None, num_classes=44).to(dev).train()
    for m in ref.modules():                       # no stochastic depth / dropout: same function twice
        if isinstance(m, torchvision.ops.StochasticDepth):
            m.p = 0.0
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    prod = copy.deepcopy(ref)
    amp = copy.deepcopy(ref).to(memory_format=torch.channels_last)       # PyTorch's own bf16 path
    assert swap_depthwise_separable(prod) == (23, 49)
    prod = prod.to(memory_format=torch.channels_last)
    x = torch.randn(16, 3, 260, 260, device=dev)         # 9x9 maps x 16 >= min_rows of the 1x1 path
    labels = torch.randint(0, 44, (16,), device=dev)
    crit = torch.nn.CrossEntropyLoss()                                  # train_advanced_v3.0.py:449
    lo = ref(x)
    loss_o = crit(lo, labels)
    loss_o.backward()
    xcl = x.contiguous(memory_format=torch.channels_last)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        la = amp(xcl)
        loss_a = crit(la.float(), labels)
    loss_a.backward()
    n0 = _lib.launch_count
    with torch.autocast("cuda", dtype=torch.bfloat16):
        lp = prod(xcl)
        loss_p = crit(lp.float(), labels)
    loss_p.backward()
    # 23 depthwise fwd + 23 x 2 bwd + 45 pointwise x (1 fwd + 2 bwd) (+ split-K zero fills are torch's)
    assert _lib.launch_count - n0 >= 23 * 3 + 45 * 3 - 1

    def gnorm(m):
        return float(torch.sqrt(sum((p.grad.double() ** 2).sum() for p in m.parameters())))

    # 100+ bf16 layers with batch-statistics BatchNorm in between: the yardstick is how far
    # PyTorch's own bf16 autocast path (cuDNN / cuBLAS kernels) lands from the fp32 model
    e_amp, e_prod = _rel(la.float(), lo), _rel(lp.float(), lo)
    assert e_prod < max(1.5 * e_amp, BF16_REL), (e_prod, e_amp)
    assert e_prod < 0.15, e_prod
    d_amp = abs(loss_a.item() - loss_o.item()) / abs(loss_o.item())
    d_prod = abs(loss_p.item() - loss_o.item()) / abs(loss_o.item())
    assert d_prod < max(1.5 * d_amp, BF16_REL), (d_prod, d_amp)
    g_o, g_a, g_p = gnorm(ref), gnorm(amp), gnorm(prod)
    assert abs(g_p - g_o) < max(1.5 * abs(g_a - g_o), 5e-2 * g_o), (g_p, g_a, g_o)
    assert all(p.grad is not None and p.grad.dtype == torch.float32 for p in prod.parameters())


# ------------------------------------------------------------------ BatchNorm + SiLU, squeeze-excite
@pytest.mark.parametrize("B,H,W,C,silu", [
    (2, 3, 5, 8, True), (1, 1, 1, 16, True), (3, 7, 9, 24, False), (2, 6, 11, 88, True),
    (1, 5, 4, 264, True), (4, 9, 9, 1408, False), (2, 33, 17, 144, True), (5, 2, 3, 2112, True)])
def test_bn_act_vs_oracle_small(cnx_lib, B, H, W, C, silu):
    """Batch-statistics BatchNorm (+SiLU) forward, running-buffer update, pooled sums and backward
    against oracle/dwsep_oracle.py (pinned to F.batch_norm / F.silu on the CPU side) at ragged
    shapes: C below / across / far above one 64-channel slab, a single pixel, odd row counts."""
    from image_classification_b200 import ops
    dev = _dev()
    rng = np.random.RandomState(B * 1000 + H * 100 + C)
    M, eps, mom = B * H * W, 1e-5, 0.1
    x = torch.from_numpy((rng.randn(B, H, W, C) * 1.5 + 0.5).astype(np.float32)).bfloat16()
    w = torch.from_numpy((rng.rand(C) + 0.5).astype(np.float32))
    b = torch.from_numpy((rng.randn(C) * 0.3).astype(np.float32))
    rm = torch.from_numpy(rng.randn(C).astype(np.float32))
    rv = torch.from_numpy((rng.rand(C) + 0.5).astype(np.float32))
    act = ops.ACT_SILU if silu else ops.ACT_NONE
    x2 = x.double().numpy().reshape(M, C)
    y_ref, mean_ref, var_ref = dwsep_oracle.bn_act_fwd_ref(x2, w.numpy(), b.numpy(), eps, silu)
    rm_ref, rv_ref = dwsep_oracle.bn_running_update_ref(rm.numpy(), rv.numpy(), mean_ref, var_ref, M, mom)
    rmd, rvd = rm.to(dev), rv.to(dev)
    y, mean, rstd, pool = ops.bn_act_fwd(x.to(dev), w.to(dev), b.to(dev), rmd, rvd, training=True,
                                         momentum=mom, eps=eps, act=act, want_pool=True)
    if M > 1:           # a single row has zero variance: y = act(bias), rstd = eps^-1/2
        assert _rel(y.cpu().reshape(M, C), torch.from_numpy(y_ref)) < 2.0 ** -7
    assert _rel(mean.cpu(), torch.from_numpy(mean_ref)) < 1e-5
    assert _rel(rstd.cpu(), torch.from_numpy(1.0 / np.sqrt(var_ref + eps))) < 1e-4
    assert _rel(rmd.cpu(), torch.from_numpy(rm_ref)) < 1e-5
    assert _rel(rvd.cpu(), torch.from_numpy(rv_ref)) < 1e-4
    pool_ref = y.float().reshape(B, H * W, C).sum(dim=1)          # sums of the ROUNDED outputs
    assert _rel(pool, pool_ref) < 1e-5
    # the same call without the pooled sums takes the flat geometry: identical y
    y2, *_ = ops.bn_act_fwd(x.to(dev), w.to(dev), b.to(dev), None, None, training=True,
                            momentum=mom, eps=eps, act=act)
    assert _rel(y2, y) < 2.0 ** -7        # statistics are atomic sums: a bf16 ulp may flip
    g = torch.from_numpy(rng.randn(B, H, W, C).astype(np.float32)).bfloat16()
    dx_ref, dw_ref, db_ref = dwsep_oracle.bn_act_bwd_ref(g.double().numpy().reshape(M, C), x2, w.numpy(),
                                                         b.numpy(), eps, silu)
    dx, dw, db = ops.bn_act_bwd(g.to(dev), x.to(dev), w.to(dev), b.to(dev), mean, rstd, act)
    if M > 1:
        assert _rel(dx.cpu().reshape(M, C), torch.from_numpy(dx_ref)) < 2.0 ** -7
        assert _rel(dw.cpu(), torch.from_numpy(dw_ref)) < 1e-4
    assert _rel(db.cpu(), torch.from_numpy(db_ref)) < 1e-4
    # eval mode: the (updated) running statistics normalise, buffers untouched
    rm_before, rv_before = rmd.clone(), rvd.clone()
    ye, *_ = ops.bn_act_fwd(x.to(dev), w.to(dev), b.to(dev), rmd, rvd, training=False,
                            momentum=mom, eps=eps, act=act)
    u = (x.double() - rm_before.cpu().double()) / torch.sqrt(rv_before.cpu().double() + eps) * w.double() + b.double()
    ye_ref = u * torch.sigmoid(u) if silu else u
    assert _rel(ye.cpu(), ye_ref) < 2.0 ** -7
    assert torch.equal(rmd, rm_before) and torch.equal(rvd, rv_before)


# (C, H) of the BatchNorm inputs of efficientnet_b2 at 260x260: expand, depthwise and project widths
EFFNET_B2_BN = [(32, 130), (16, 130), (96, 130), (96, 65), (24, 65), (144, 65), (144, 33), (48, 33),
                (288, 33), (288, 17), (88, 17), (528, 17), (120, 17), (720, 17), (720, 9), (208, 9),
                (1248, 9), (352, 9), (2112, 9), (1408, 9)]


@pytest.mark.parametrize("C,H", EFFNET_B2_BN)
def test_bn_silu_effnet_b2_shapes_vs_torch_fp64(cnx_lib, C, H):
    """Batch 48 (previous/V3-efb2/config.py:14) at the network's own shapes against
    F.batch_norm(training) -> F.silu and autograd in float64 on the device."""
    from image_classification_b200 import ops
    dev = _dev()
    B = 48
    torch.manual_seed(C * 7 + H)
    x = (torch.randn(B, H, H, C, device=dev) * 1.3 + 0.2).bfloat16()
    w = torch.rand(C, device=dev) + 0.5
    b = torch.randn(C, device=dev) * 0.2
    xr = x.double().reshape(-1, C).requires_grad_()
    wr, br = w.double().requires_grad_(), b.double().requires_grad_()
    yr = F.silu(F.batch_norm(xr, None, None, wr, br, True, 0.1, 1e-5))
    g = torch.randn(B, H, H, C, device=dev).bfloat16()
    yr.backward(g.double().reshape(-1, C))
    y, mean, rstd, _ = ops.bn_act_fwd(x, w, b, None, None, training=True, momentum=0.1, eps=1e-5,
                                      act=ops.ACT_SILU)
    assert _rel(y.reshape(-1, C), yr.detach()) < 2.0 ** -7                # bf16 rounding of y
    dx, dw, db = ops.bn_act_bwd(g, x, w, b, mean, rstd, ops.ACT_SILU)
    assert _rel(dx.reshape(-1, C), xr.grad) < BF16_REL
    assert _rel(dw, wr.grad) < 1e-3                                       # fp32 sums over ~10^5..10^6 rows
    assert _rel(db, br.grad) < 1e-3


@pytest.mark.parametrize("B,H,W,C,Cs", [(2, 3, 5, 8, 2), (3, 1, 1, 16, 4), (2, 9, 7, 96, 4),
                                        (4, 17, 17, 528, 22), (2, 5, 6, 2112, 88)])
def test_squeeze_excite_vs_oracle(cnx_lib, B, H, W, C, Cs):
    """se_scale kernels + SqueezeExciteB200 (fc1 / fc2 as fp32 [B,C] layers) against the oracle
    (pinned to torchvision SqueezeExcitation): output, input gradient, all four parameter grads;
    with and without the pooled sums handed over by the preceding BatchNorm pass."""
    from image_classification_b200.efficientnet import SqueezeExciteB200
    dev = _dev()
    torch.manual_seed(B * 100 + C)
    fc1 = torch.nn.Conv2d(C, Cs, 1).to(dev)
    fc2 = torch.nn.Conv2d(Cs, C, 1).to(dev)
    with torch.no_grad():
        fc1.weight.mul_(3.0), fc2.weight.mul_(3.0)            # gates away from 0.5
    se = SqueezeExciteB200(fc1, fc2)
    x = torch.randn(B, C, H, W, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    g = torch.randn(B, C, H, W, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    w1, b1 = fc1.weight.detach().reshape(Cs, C).cpu().numpy(), fc1.bias.detach().cpu().numpy()
    w2, b2 = fc2.weight.detach().reshape(C, Cs).cpu().numpy(), fc2.bias.detach().cpu().numpy()
    xn = x.permute(0, 2, 3, 1).reshape(B, H * W, C).double().cpu().numpy()
    gn = g.permute(0, 2, 3, 1).reshape(B, H * W, C).double().cpu().numpy()
    y_ref, _ = dwsep_oracle.se_fwd_ref(xn, w1, b1, w2, b2)
    dx_ref, dw1, db1, dw2, db2 = dwsep_oracle.se_bwd_ref(gn, xn, w1, b1, w2, b2)
    for handed in (False, True):
        # without the hand-over the pooled mean comes from cnx_avgpool_nhwc_fwd as bf16 (what
        # PyTorch's bf16 adaptive_avg_pool2d returns too): parameter gradients carry that rounding
        ptol = 1e-3 if handed else 6e-3
        xa = x.clone().requires_grad_()
        if handed:
            xa._cnx_pool_sum = x.permute(0, 2, 3, 1).float().sum(dim=(1, 2))
        for p in se.parameters():
            p.grad = None
        y = se(xa)
        y.backward(g)
        flat = lambda t: t.permute(0, 2, 3, 1).reshape(B, H * W, C).cpu()
        assert _rel(flat(y), torch.from_numpy(y_ref)) < 2.0 ** -7
        assert _rel(flat(xa.grad), torch.from_numpy(dx_ref)) < 2.0 ** -7
        assert _rel(fc1.weight.grad.reshape(Cs, C).cpu(), torch.from_numpy(dw1)) < ptol
        assert _rel(fc1.bias.grad.cpu(), torch.from_numpy(db1)) < ptol
        assert _rel(fc2.weight.grad.reshape(C, Cs).cpu(), torch.from_numpy(dw2)) < ptol
        assert _rel(fc2.bias.grad.cpu(), torch.from_numpy(db2)) < ptol


def test_bn_se_reject_bad_arguments(cnx_lib):
    from image_classification_b200 import ops, _lib
    dev = _dev()
    x = torch.zeros(2, 4, 4, 16, device=dev, dtype=torch.bfloat16)
    w = torch.ones(16, device=dev)
    with pytest.raises(_lib.CnxError):                        # C % 8
        ops.bn_act_fwd(x[..., :12].contiguous(), w[:12], w[:12], None, None, training=True,
                       momentum=0.1, eps=1e-5, act=0)
    with pytest.raises(_lib.CnxError):                        # unknown activation
        ops.bn_act_fwd(x, w, w, None, None, training=True, momentum=0.1, eps=1e-5, act=7)
    with pytest.raises(_lib.CnxError):                        # eval mode without running statistics
        ops.bn_act_fwd(x, w, w, None, None, training=False, momentum=0.1, eps=1e-5, act=0)
    with pytest.raises(_lib.CnxError):                        # fp32 activations
        ops.bn_act_fwd(x.float(), w, w, None, None, training=True, momentum=0.1, eps=1e-5, act=0)
    with pytest.raises(_lib.CnxError):                        # gate of the wrong shape
        ops.se_scale_fwd(x, torch.ones(2, 8, device=dev))


def test_efficientnet_b2_fully_swapped_vs_fp32_model(cnx_lib):
    """torchvision efficientnet_b2 with convolutions, BatchNorm+SiLU and squeeze-excite all on
    libcnx_b200 (swap_efficientnet) under bf16 autocast, one training forward/backward with the
    reference's criterion, against the untouched fp32 model; PyTorch's own bf16 autocast path is
    the yardstick for how far 100+ bf16 layers may drift.  Running statistics must match too."""
    import copy
    import torchvision
    from image_classification_b200 import _lib
    from image_classification_b200.efficientnet import swap_efficientnet
    dev = _dev()
    torch.manual_seed(0)
    ref = torchvision.models.efficientnet_b2(weights=None, num_classes=44).to(dev).train()
    for m in ref.modules():
        if isinstance(m, (torchvision.ops.StochasticDepth, torch.nn.Dropout)):
            m.p = 0.0
    prod = copy.deepcopy(ref)
    amp = copy.deepcopy(ref).to(memory_format=torch.channels_last)
    counts = swap_efficientnet(prod)
    assert counts == {"depthwise": 23, "pointwise": 49, "batchnorm": 69, "squeeze_excite": 23}
    prod = prod.to(memory_format=torch.channels_last)
    x = torch.randn(16, 3, 260, 260, device=dev)
    labels = torch.randint(0, 44, (16,), device=dev)
    crit = torch.nn.CrossEntropyLoss()                                  # train_advanced_v3.0.py:449
    lo = ref(x)
    loss_o = crit(lo, labels)
    loss_o.backward()
    xcl = x.contiguous(memory_format=torch.channels_last)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        la = amp(xcl)
        loss_a = crit(la.float(), labels)
    loss_a.backward()
    n0 = _lib.launch_count
    with torch.autocast("cuda", dtype=torch.bfloat16):
        lp = prod(xcl)
        loss_p = crit(lp.float(), labels)
    loss_p.backward()
    # + 68 BatchNorm x (stats + apply + 2 bwd) + 23 squeeze-excite x (scale + reduce + apply)
    assert _lib.launch_count - n0 >= 23 * 3 + 45 * 3 + 68 * 4 + 23 * 3 - 1

    def gnorm(m):
        return float(torch.sqrt(sum((p.grad.double() ** 2).sum() for p in m.parameters())))

    e_amp, e_prod = _rel(la.float(), lo), _rel(lp.float(), lo)
    assert e_prod < max(1.5 * e_amp, BF16_REL), (e_prod, e_amp)
    d_amp = abs(loss_a.item() - loss_o.item()) / abs(loss_o.item())
    d_prod = abs(loss_p.item() - loss_o.item()) / abs(loss_o.item())
    assert d_prod < max(1.5 * d_amp, BF16_REL), (d_prod, d_amp)
    g_o, g_a, g_p = gnorm(ref), gnorm(amp), gnorm(prod)
    assert abs(g_p - g_o) < max(1.5 * abs(g_a - g_o), 5e-2 * g_o), (g_p, g_a, g_o)
    assert all(p.grad is not None and p.grad.dtype == torch.float32 for p in prod.parameters())
    # running statistics after one update (momentum 0.1): variances relative to their size, means
    # against the layer's standard deviation (the expand convolutions see zero-mean inputs, so
    # their batch means are ~1e-7 of the std: a relative measure of the mean itself is noise)
    pm, rm = dict(prod.named_modules()), dict(ref.named_modules())
    n_bn = 0
    for name, bn in rm.items():
        if not isinstance(bn, torch.nn.BatchNorm2d):
            continue
        n_bn += 1
        assert int(pm[name].num_batches_tracked) == int(bn.num_batches_tracked) == 1
        std = float(bn.running_var.sqrt().max())
        assert float((pm[name].running_mean - bn.running_mean).abs().max()) < 2e-2 * std, name
        assert _rel(pm[name].running_var, bn.running_var) < 5e-2, name
    assert n_bn == 69


def test_graphed_forward_backward_matches_eager(cnx_lib):
    """graphed.GraphedForwardBackward on swapped efficientnet_b2 stages: the replayed
    micro-step gives the eager loss / logits / gradients for each of two different batches (the
    static input is refilled, the static gradients are overwritten, not accumulated) and advances
    the BatchNorm buffers once per replay."""
    import copy
    import torchvision
    from image_classification_b200.efficientnet import swap_efficientnet
    from image_classification_b200.graphed import GraphedForwardBackward
    from image_classification_b200.loss import SoftTargetCrossEntropyB200
    dev = _dev()
    torch.manual_seed(1)
    full = torchvision.models.efficientnet_b2(weights=None, num_classes=44)
    for m in full.modules():
        if isinstance(m, (torchvision.ops.StochasticDepth, torch.nn.Dropout)):
            m.p = 0.0
    # stem + the first three MBConv stages (8 blocks, every kind of library module) and a head: the
    # full 23-block network at random init amplifies the fp32 summation-order noise of the
    # BatchNorm statistics (atomics) to several percent of the logits, eager against eager
    base = torch.nn.Sequential(full.features[:4], torch.nn.AdaptiveAvgPool2d(1), torch.nn.Flatten(),
                               torch.nn.Linear(48, 44)).to(dev).train()
    counts = swap_efficientnet(base)
    assert counts["batchnorm"] >= 20 and counts["squeeze_excite"] == 8
    eager, graphed = copy.deepcopy(base), copy.deepcopy(base)
    crit = SoftTargetCrossEntropyB200(0.0)
    scaler = torch.amp.GradScaler("cuda", init_scale=1024.0)
    batches = [(torch.randn(8, 3, 160, 160, device=dev).contiguous(memory_format=torch.channels_last),
                torch.randint(0, 44, (8,), device=dev)) for _ in range(2)]
    warm = 3
    gfb = GraphedForwardBackward(graphed, crit, scaler, *batches[0], warmup=warm)
    bn_e = [m for m in eager.modules() if isinstance(m, torch.nn.BatchNorm2d)][0]
    bn_g = [m for m in graphed.modules() if isinstance(m, torch.nn.BatchNorm2d)][0]
    assert int(bn_g.num_batches_tracked) == warm              # capture records, it does not run
    for _ in range(warm):                                     # bring the eager buffers level
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            eager(batches[0][0])
    for i, (x, y) in enumerate(batches):
        for p in eager.parameters():
            p.grad = None
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = eager(x)
            loss = crit(out, y)
        scaler.scale(loss).backward()
        loss_g = gfb(x, y)
        assert abs(float(loss_g) - float(loss)) < BF16_REL * abs(float(loss))
        assert _rel(gfb.outputs.float(), out.float()) < BF16_REL
        ge = torch.cat([p.grad.flatten() for p in eager.parameters()])
        gg = torch.cat([p.grad.flatten() for p in graphed.parameters()])
        assert float((ge - gg).norm() / ge.norm()) < BF16_REL, i      # atomics reorder fp32 sums
        assert int(bn_g.num_batches_tracked) == int(bn_e.num_batches_tracked) == warm + 1 + i
        assert _rel(bn_g.running_var, bn_e.running_var) < 1e-3
    with pytest.raises(RuntimeError):
        gfb(batches[0][0][:4], batches[0][1][:4])             # captured for a fixed batch shape


def test_timm_layout_efficientnet_swapped_vs_fp32(cnx_lib):
    """The timm-layout stand-in (tests/_foreign.py: BatchNormAct2d, conv_reduce / conv_expand
    squeeze-excite, DepthwiseSeparableConv / InvertedResidual) fully swapped under bf16 autocast,
    one training forward/backward against the untouched fp32 model: logits, loss, every parameter
    gradient as one vector, running statistics."""
    import copy
    from _foreign import TimmLikeEfficientNet
    from image_classification_b200 import _lib
    from image_classification_b200.efficientnet import swap_efficientnet
    dev = _dev()
    torch.manual_seed(2)
    ref = TimmLikeEfficientNet().to(dev).train()
    prod = copy.deepcopy(ref)
    assert swap_efficientnet(prod)["batchnorm"] == 18
    prod = prod.to(memory_format=torch.channels_last)
    x = torch.randn(16, 3, 192, 192, device=dev)
    labels = torch.randint(0, 44, (16,), device=dev)
    crit = torch.nn.CrossEntropyLoss()
    lo = ref(x)
    loss_o = crit(lo, labels)
    loss_o.backward()
    n0 = _lib.launch_count
    with torch.autocast("cuda", dtype=torch.bfloat16):
        lp = prod(x.contiguous(memory_format=torch.channels_last))
        loss_p = crit(lp.float(), labels)
    loss_p.backward()
    # 6 depthwise x 3 + 18 BatchNorm x 4 + 6 squeeze-excite x (2 + 1 + 1 + 3 + 1) at least
    assert _lib.launch_count - n0 >= 6 * 3 + 18 * 4 + 6 * 8
    assert _rel(lp.float(), lo) < 5e-2                       # 18 bf16 layers with batch statistics
    assert abs(loss_p.item() - loss_o.item()) < BF16_REL * abs(loss_o.item())
    go = torch.cat([p.grad.flatten() for p in ref.parameters()])
    gp = torch.cat([p.grad.flatten() for p in prod.parameters()])
    assert float((go - gp).norm() / go.norm()) < 5e-2
    pm = dict(prod.named_modules())
    for name, bn in ref.named_modules():
        if isinstance(bn, torch.nn.BatchNorm2d):
            std = float(bn.running_var.sqrt().max())
            assert float((pm[name].running_mean - bn.running_mean).abs().max()) < 2e-2 * std, name
            assert _rel(pm[name].running_var, bn.running_var) < 5e-2, name


def test_bn_statistics_of_channels_far_from_zero(cnx_lib):
    """Channel means 60 standard deviations from zero over 10^5 rows: the statistics are shifted
    sums (x - x[0,c]), so E[d^2] - E[d]^2 does not cancel; mean, rstd and the normalised output
    stay at the fp32 / bf16 level of the float64 oracle."""
    from image_classification_b200 import ops
    dev = _dev()
    torch.manual_seed(11)
    B, H, W, C = 32, 56, 56, 24
    offs = torch.linspace(-60.0, 60.0, C, device=dev)
    x = (offs + 0.5 * torch.randn(B, H, W, C, device=dev)).bfloat16()
    w = torch.rand(C, device=dev) + 0.5
    b = torch.randn(C, device=dev) * 0.1
    y, mean, rstd, _ = ops.bn_act_fwd(x, w, b, None, None, training=True, momentum=0.1, eps=1e-5,
                                      act=ops.ACT_NONE)
    y_ref, mean_ref, var_ref = dwsep_oracle.bn_act_fwd_ref(
        x.double().cpu().numpy().reshape(-1, C), w.cpu().numpy(), b.cpu().numpy(), 1e-5, False)
    assert _rel(mean.cpu(), torch.from_numpy(mean_ref)) < 1e-6
    assert float((rstd.cpu().double() * torch.from_numpy(np.sqrt(var_ref + 1e-5)) - 1).abs().max()) < 1e-4
    assert _rel(y.cpu().reshape(-1, C), torch.from_numpy(y_ref)) < 2.0 ** -7


def test_averaged_model_matches_torch_swa(cnx_lib):
    """ema.AveragedModelB200 vs torch.optim.swa_utils.AveragedModel (the reference's SWA,
    previous/V3-efb2/train_advanced_v3.0.py:445, :505-507) over four updates of a drifting model:
    same state_dict keys, parameters equal to fp32 rounding, n_averaged counts."""
    import copy
    from torch.optim.swa_utils import AveragedModel
    from _foreign import TimmLikeEfficientNet
    from image_classification_b200.ema import AveragedModelB200
    dev = _dev()
    torch.manual_seed(4)
    model = TimmLikeEfficientNet().to(dev)
    ref, swa = AveragedModel(copy.deepcopy(model)), AveragedModelB200(model)
    assert list(swa.state_dict()) == list(ref.state_dict())
    for step in range(4):
        with torch.no_grad():
            for p in model.parameters():
                p.add_(torch.randn_like(p) * 0.05)
        ref.update_parameters(model)
        swa.update_parameters(model)
        assert int(swa.n_averaged) == int(ref.n_averaged) == step + 1
        for (n, a), (_, b) in zip(swa.module.named_parameters(), ref.module.named_parameters()):
            assert torch.allclose(a, b, rtol=1e-5, atol=1e-7), (step, n)
    x = torch.randn(2, 3, 64, 64, device=dev)
    swa.eval(), ref.eval()
    assert torch.allclose(swa(x), ref(x), rtol=1e-4, atol=1e-5)
