"""GPU parity (through the C ABI) of the ConvNeXt block kernels against the oracle.

Tolerances (BASELINE.json north_star): bf16 tensors within 2e-2 relative, fp32 statistics within
1e-4 relative.  H1: layer-scale 1e-6 hides the branch, so block-level checks use
gamma ~ U(0.1, 1) AND gamma = 1e-6, compare the BRANCH (y - x) and every parameter gradient.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from oracle import block_oracle  # noqa: E402  (checker only)

BF16_REL = 2e-2
FP32_REL = 1e-4


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a B200"
    return torch.device("cuda:0")


def _rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _bf(t):
    """round to bf16, keep fp32 storage (what the kernels actually see)"""
    return t.bfloat16().float()


# ---------------------------------------------------------------------------- dwconv + LN fwd
DW_SHAPES = [(2, 7, 7, 64), (1, 14, 14, 128), (2, 9, 11, 96), (1, 12, 12, 256), (2, 14, 14, 512),
             (1, 7, 7, 1024), (1, 56, 56, 128), (1, 24, 20, 192), (2, 5, 3, 384), (1, 28, 28, 256),
             (1, 1, 1, 64), (1, 40, 36, 128)]


@pytest.mark.parametrize("shape", DW_SHAPES)
def test_dwconv_ln_fwd_vs_oracle(cnx_lib, shape):
    from image_classification_b200 import ops
    B, H, W, C = shape
    dev = _dev()
    torch.manual_seed(C + H)
    x = _bf(torch.randn(B, H, W, C, device=dev) * 1.5 + 0.3)
    w = torch.randn(C, 1, 7, 7, device=dev) * 0.15
    b = torch.randn(C, device=dev) * 0.5
    ref_xhat, ref_mean, ref_rstd = block_oracle.dwconv_ln_ref(x.double(), w.double(), b.double())
    # statistics are taken over the bf16-rounded conv output, as under the reference's autocast
    _, amp_mean, amp_rstd = block_oracle.dwconv_ln_ref(x.double(), w.double(), b.double(),
                                                       amp_round=torch.bfloat16)
    wt = w.reshape(C, 49).t().contiguous()
    xhat, mean, rstd = ops.dwconv_ln_fwd(x.bfloat16(), wt, b)
    assert xhat.shape == (B, H, W, C) and xhat.dtype == torch.bfloat16
    # fp32 kernels vs the autocast-numerics oracle: 1e-4 (a conv sum within half a bf16 ulp of a
    # rounding boundary may round the other way than fp64 does: allow it in the mean)
    assert _rel(mean, amp_mean) < FP32_REL or (mean.double() - amp_mean).abs().max() < 2.0 ** -8 * 8 / C
    assert _rel(rstd, amp_rstd) < FP32_REL + 0.1 / C
    # and both stay within bf16 tolerance of the exact (un-rounded) statistics
    assert _rel(mean, ref_mean) < BF16_REL and _rel(rstd, ref_rstd) < BF16_REL
    assert _rel(xhat, ref_xhat) < BF16_REL
    # tighter: two roundings are left — the conv output is rounded to bf16 before LayerNorm (as
    # under the reference's autocast, where conv2d returns half precision) and the final store
    assert (xhat.double() - ref_xhat).abs().max() < 2.0 ** -7 * ref_xhat.abs().max() + 1e-3


def test_dwconv_ln_fwd_rejects_bad_shapes(cnx_lib):
    from image_classification_b200 import _lib, ops
    dev = _dev()
    with pytest.raises(_lib.CnxError):      # C not a multiple of 8
        ops.dwconv_ln_fwd(torch.zeros(1, 4, 4, 36, device=dev, dtype=torch.bfloat16),
                          torch.zeros(49, 36, device=dev), torch.zeros(36, device=dev))
    with pytest.raises(_lib.CnxError):      # fp32 activations are not accepted silently
        ops.dwconv_ln_fwd(torch.zeros(1, 4, 4, 64, device=dev),
                          torch.zeros(49, 64, device=dev), torch.zeros(64, device=dev))


# ------------------------------------------------------------------------------------ GEMM
GEMM_SHAPES = [(128, 128, 64), (256, 384, 128), (200, 136, 72), (98, 96, 384), (1000, 512, 128),
               (6272, 1024, 256), (129, 8, 8), (25088, 256, 64)]


@pytest.mark.parametrize("shape", GEMM_SHAPES)
def test_gemm_plain_k_major(cnx_lib, shape):
    from image_classification_b200 import ops
    M, N, K = shape
    dev = _dev()
    torch.manual_seed(M + N + K)
    A = torch.randn(M, K, device=dev).bfloat16()
    Bm = torch.randn(N, K, device=dev).bfloat16()
    ref = A.double() @ Bm.double().t()
    D = ops.gemm_bf16(A, Bm)
    assert _rel(D, ref) < BF16_REL
    assert (D.double() - ref).abs().max() < 2.0 ** -8 * ref.abs().max() + 1e-2
    D32 = ops.gemm_bf16(A, Bm, epilogue=ops.EPI_STORE_F32)      # fp32 out (split-K when small)
    assert _rel(D32, ref) < 1e-4


@pytest.mark.parametrize("a_mn,b_mn", [(True, True), (True, False), (False, True)])
@pytest.mark.parametrize("shape", [(128, 128, 64), (96, 384, 98), (512, 128, 6272), (136, 200, 1000)])
def test_gemm_mn_major_operands(cnx_lib, shape, a_mn, b_mn):
    """Weight-gradient form: operands stored [K, rows]; K (tokens) may be ragged."""
    from image_classification_b200 import ops
    M, N, K = shape
    if K % 8 != 0 and not (a_mn and b_mn):
        pytest.skip("a K-major operand needs K % 8 == 0 (16-byte TMA row pitch)")
    dev = _dev()
    torch.manual_seed(M * 3 + N + K)
    A = torch.randn(M, K, device=dev).bfloat16()
    Bm = torch.randn(N, K, device=dev).bfloat16()
    ref = A.double() @ Bm.double().t()
    Ain = A.t().contiguous() if a_mn else A
    Bin = Bm.t().contiguous() if b_mn else Bm
    D = ops.gemm_bf16(Ain, Bin, a_mn_major=a_mn, b_mn_major=b_mn, epilogue=ops.EPI_STORE_F32)
    assert _rel(D, ref) < 1e-4


def test_gemm_epilogues(cnx_lib):
    from image_classification_b200 import ops
    dev = _dev()
    torch.manual_seed(5)
    M, N, K = 300, 264, 136
    A = (torch.randn(M, K, device=dev) * 0.3).bfloat16()
    Bm = (torch.randn(N, K, device=dev) * 0.3).bfloat16()
    bias = torch.randn(N, device=dev)
    R = torch.randn(M, N, device=dev).bfloat16()
    acc = A.double() @ Bm.double().t()
    # fc1: z = acc + bias (bf16), h = gelu(z_bf16)
    z, h = ops.gemm_bf16(A, Bm, epilogue=ops.EPI_BIAS_GELU, bias=bias)
    ref_z = acc + bias.double()
    assert _rel(z, ref_z) < BF16_REL
    ref_h = F.gelu(z.double())
    assert _rel(h, ref_h) < BF16_REL
    assert (h.double() - ref_h).abs().max() < 2.0 ** -8 * ref_h.abs().max() + 1e-3
    # fc2: y = acc + bias + R
    y = ops.gemm_bf16(A, Bm, epilogue=ops.EPI_BIAS_RESIDUAL, bias=bias, R=R)
    assert _rel(y, acc + bias.double() + R.double()) < BF16_REL
    # dz = acc * gelu'(R), colsum
    colsum = torch.zeros(N, device=dev)
    dz = ops.gemm_bf16(A, Bm, epilogue=ops.EPI_GELU_GRAD, R=R, colsum=colsum)
    Rd = R.double().requires_grad_()
    F.gelu(Rd).sum().backward()
    ref_dz = acc * Rd.grad
    assert _rel(dz, ref_dz) < BF16_REL
    assert _rel(colsum, dz.double().sum(0)) < 1e-4      # sums of the rounded dz


@pytest.mark.parametrize("epi_warps", [8, 16, 64])
@pytest.mark.parametrize("tuning", [(2, 0, 256, -1), (-1, 0, 128, -1), (-1, 1, 256, -1), (3, 1, 128, 3),
                                    (2, 1, 256, 5)])
def test_gemm_forced_tilings(cnx_lib, tuning, epi_warps):
    """Every kernel instantiation (single CTA / CTA pair, 128- and 256-wide tiles, shallow operand
    ring, forced split-K, 8 or 16 epilogue warps, 32- or 64-column staging chunks) gives the same
    results: ragged M, N and K, several tiles per CTA.  epi_warps == 64 stands for 8 warps with
    64-column staging chunks forced (8 alone: 32-column chunks where the epilogue has no residual)."""
    from image_classification_b200 import ops, _lib
    dev = _dev()
    torch.manual_seed(11)
    M, N, K = 1000, 520, 200
    A = (torch.randn(M, K, device=dev) * 0.3).bfloat16()
    Bm = (torch.randn(N, K, device=dev) * 0.3).bfloat16()
    bias = torch.randn(N, device=dev)
    R = torch.randn(M, N, device=dev).bfloat16()
    acc = A.double() @ Bm.double().t()
    lib = _lib.load()
    lib.cnx_debug_gemm_tuning(*tuning)
    lib.cnx_debug_gemm_epilogue_warps(8 if epi_warps == 64 else epi_warps)
    lib.cnx_debug_gemm_staging_chunk(64 if epi_warps == 64 else 0)
    try:
        D = ops.gemm_bf16(A, Bm)
        z, h = ops.gemm_bf16(A, Bm, epilogue=ops.EPI_BIAS_GELU, bias=bias)
        y = ops.gemm_bf16(A, Bm, epilogue=ops.EPI_BIAS_RESIDUAL, bias=bias, R=R)
        colsum = torch.zeros(N, device=dev)
        dz = ops.gemm_bf16(A, Bm, epilogue=ops.EPI_GELU_GRAD, R=R, colsum=colsum)
        D32 = ops.gemm_bf16(A, Bm, epilogue=ops.EPI_STORE_F32)
        Dw = ops.gemm_bf16(A.t().contiguous(), Bm.t().contiguous(), a_mn_major=True,
                           b_mn_major=True, epilogue=ops.EPI_STORE_F32)
        torch.cuda.synchronize()
    finally:
        lib.cnx_debug_gemm_tuning(-1, -1, -1, -1)
        lib.cnx_debug_gemm_epilogue_warps(0)
        lib.cnx_debug_gemm_staging_chunk(0)
    assert lib.cnx_debug_timeout_flag() == 0
    assert _rel(D, acc) < BF16_REL
    assert _rel(z, acc + bias.double()) < BF16_REL
    assert _rel(h, F.gelu(z.double())) < BF16_REL
    assert _rel(y, acc + bias.double() + R.double()) < BF16_REL
    Rd = R.double().requires_grad_()
    F.gelu(Rd).sum().backward()
    assert _rel(dz, acc * Rd.grad) < BF16_REL
    assert _rel(colsum, dz.double().sum(0)) < 1e-4
    assert _rel(D32, acc) < 1e-4
    assert _rel(Dw, acc) < 1e-4


# ---------------------------------------------------------------------------- dwconv + LN bwd
@pytest.mark.parametrize("shape", [(2, 7, 7, 64), (1, 14, 14, 128), (2, 9, 11, 96), (1, 12, 12, 256),
                                   (2, 14, 14, 512), (1, 7, 7, 1024), (1, 28, 28, 128),
                                   (1, 24, 20, 192), (1, 56, 56, 128)])
def test_dwconv_ln_bwd_vs_oracle_autograd(cnx_lib, shape):
    from image_classification_b200 import ops
    B, H, W, C = shape
    dev = _dev()
    torch.manual_seed(C + W)
    x = _bf(torch.randn(B, H, W, C, device=dev)).double().requires_grad_()
    w = (torch.randn(C, 1, 7, 7, device=dev) * 0.15).double().requires_grad_()
    b = (torch.randn(C, device=dev) * 0.5).double().requires_grad_()
    dxhat = _bf(torch.randn(B, H, W, C, device=dev))
    dy = _bf(torch.randn(B, H, W, C, device=dev))
    xhat_ref, _, rstd_ref = block_oracle.dwconv_ln_ref(x, w, b)
    (xhat_ref * dxhat.double()).sum().backward()
    ref_dx = x.grad + dy.double()
    ref_dwt = w.grad.reshape(C, 49).t()
    # kernel inputs: what the forward kernel would have saved
    xhat_in = xhat_ref.detach().float().bfloat16().contiguous()
    rstd_in = rstd_ref.detach().float().contiguous()
    wt = w.detach().float().reshape(C, 49).t().contiguous()
    dx, dwt, dbias, dycs = ops.dwconv_ln_bwd(dxhat.bfloat16(), xhat_in, rstd_in,
                                             x.detach().float().bfloat16(), dy.bfloat16(), wt)
    assert _rel(dx, ref_dx) < BF16_REL
    assert _rel(dwt, ref_dwt) < BF16_REL
    assert _rel(dbias, b.grad) < BF16_REL
    assert _rel(dycs, dy.double().sum((0, 1, 2))) < 1e-4


# ------------------------------------------------------------------------------------ block
def _make_pair(C, gamma_mode, dev):
    from image_classification_b200.block import ConvNeXtBlockB200
    torch.manual_seed(C)
    ref = block_oracle.OracleConvNeXtBlock(C).to(dev)
    with torch.no_grad():
        for n, p in ref.named_parameters():
            if n == "gamma":
                p.copy_(torch.rand_like(p) * 0.9 + 0.1 if gamma_mode == "o1" else
                        torch.full_like(p, 1e-6))
            elif n == "norm.weight":
                p.copy_(1 + 0.2 * torch.randn_like(p))
            elif p.dim() > 1:
                p.copy_(torch.randn_like(p) * (0.15 if "conv" in n else 0.06))
            else:
                p.copy_(torch.randn_like(p) * 0.2)
    blk = ConvNeXtBlockB200(C).to(dev).load_from(ref)
    return ref, blk


@pytest.mark.parametrize("gamma_mode", ["o1", "1e-6"])
@pytest.mark.parametrize("shape", [(2, 64, 7, 7), (2, 96, 14, 14), (1, 128, 28, 28), (2, 256, 14, 14),
                                   (1, 512, 14, 14), (1, 1024, 7, 7)])
def test_block_forward_backward_vs_oracle(cnx_lib, shape, gamma_mode):
    B, C, H, W = shape
    dev = _dev()
    ref, blk = _make_pair(C, gamma_mode, dev)
    torch.manual_seed(1)
    x32 = _bf(torch.randn(B, C, H, W, device=dev))
    gy = _bf(torch.randn(B, C, H, W, device=dev))
    # oracle in fp64 on the same (bf16-representable) inputs
    ref = ref.double()
    xr = x32.double().requires_grad_()
    yr = ref(xr)
    yr.backward(gy.double())
    # product: bf16 channels_last fast path
    xb = x32.bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_()
    yb = blk(xb)
    assert yb.dtype == torch.bfloat16 and yb.shape == xb.shape
    yb.backward(gy.bfloat16().contiguous(memory_format=torch.channels_last))
    if gamma_mode == "o1":
        branch, branch_ref = yb.double() - x32.double(), (yr - xr).detach()
        assert _rel(branch, branch_ref) < BF16_REL + 2.0 ** -8 * float(yr.abs().max() / branch_ref.abs().max())
    assert _rel(yb, yr.detach()) < BF16_REL
    assert _rel(xb.grad, xr.grad) < BF16_REL
    for (n, p), (_, q) in zip(blk.named_parameters(), ref.named_parameters()):
        assert p.grad is not None and p.grad.dtype == torch.float32, n
        assert _rel(p.grad, q.grad) < BF16_REL, (n, _rel(p.grad, q.grad))


def test_block_fp32_nchw_drop_in_path(cnx_lib):
    """What a timm stage under the reference's autocast hands the block: fp32 contiguous NCHW."""
    dev = _dev()
    ref, blk = _make_pair(128, "o1", dev)
    torch.manual_seed(2)
    x = _bf(torch.randn(2, 128, 14, 14, device=dev)).requires_grad_()
    y = blk(x)
    assert y.dtype == torch.float32 and y.is_contiguous()
    y.backward(torch.ones_like(y))
    xr = x.detach().double().requires_grad_()
    yr = ref.double()(xr)
    yr.backward(torch.ones_like(yr))
    assert _rel(y, yr.detach()) < BF16_REL
    assert _rel(x.grad, xr.grad) < BF16_REL


def test_block_state_dict_is_timm_named(cnx_lib):
    from image_classification_b200.block import ConvNeXtBlockB200
    blk = ConvNeXtBlockB200(64)
    assert sorted(blk.state_dict()) == sorted([
        "conv_dw.weight", "conv_dw.bias", "norm.weight", "norm.bias", "mlp.fc1.weight",
        "mlp.fc1.bias", "mlp.fc2.weight", "mlp.fc2.bias", "gamma"])
    assert blk.conv_dw.out_channels == 64 and tuple(blk.conv_dw.weight.shape) == (64, 1, 7, 7)
    assert float(blk.gamma[0]) == pytest.approx(1e-6)
    from image_classification_b200 import _lib
    with pytest.raises(_lib.CnxError):
        blk(torch.zeros(1, 64, 7, 7))         # CPU tensor: loud failure, no fallback


def test_weight_fold_cache_tracks_versions(cnx_lib):
    dev = _dev()
    _, blk = _make_pair(64, "o1", dev)
    x = torch.randn(1, 64, 7, 7, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    y0 = blk(x)
    f0 = blk._fold
    assert blk(x) is not None and blk._fold is f0            # same weights -> cached
    with torch.no_grad():
        blk.mlp.fc2.bias.add_(1.0)                            # in-place update bumps _version
    y1 = blk(x)
    assert blk._fold is not f0
    assert not torch.equal(y0, y1)


def test_block_sub_batched_path_equals_single_pass(cnx_lib, monkeypatch):
    """CNX_L2_CHUNK_MB (block._l2_sub_batches, off by default): the block run sub-batch by sub-batch
    through all its kernels gives the same output, input gradient and parameter gradients."""
    from image_classification_b200 import block as block_mod
    from image_classification_b200.block import ConvNeXtBlockB200
    dev = _dev()
    torch.manual_seed(4)
    blk = ConvNeXtBlockB200(128).to(dev)
    with torch.no_grad():
        blk.gamma.copy_(torch.rand(128, device=dev) * 0.9 + 0.1)
    x = torch.randn(8, 128, 12, 10, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    gy = torch.randn_like(x)
    outs = []
    for mb in ("0", "0.3"):                 # hidden tensor: 8*120*512*2 B = 0.98 MB -> 4 sub-batches
        monkeypatch.setenv("CNX_L2_CHUNK_MB", mb)
        assert block_mod._l2_sub_batches(8, 12, 10, 128) == (1 if mb == "0" else 4)
        blk.zero_grad(set_to_none=True)
        xi = x.clone().requires_grad_()
        y = blk(xi)
        y.backward(gy)
        outs.append((y.detach(), xi.grad, {n: p.grad.clone() for n, p in blk.named_parameters()}))
    (y0, dx0, g0), (y1, dx1, g1) = outs
    assert torch.equal(y0, y1) and torch.equal(dx0, dx1)        # per-sample arithmetic is the same
    for n in g0:
        assert _rel(g1[n], g0[n]) < 1e-4, n                      # sums in a different order


def test_block_backward_ln_by_linearity_route(cnx_lib, monkeypatch):
    """CNX_LNBWD_LINEARITY=1 (block.py): the block's backward with the LayerNorm backward taken in
    the GEMM epilogues gives the row-kernel route's gradients within bf16 rounding."""
    from image_classification_b200.block import ConvNeXtBlockB200
    dev = _dev()
    torch.manual_seed(6)
    blk = ConvNeXtBlockB200(256).to(dev)
    with torch.no_grad():
        blk.gamma.copy_(torch.rand(256, device=dev) * 0.9 + 0.1)
    x = torch.randn(4, 256, 14, 14, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    gy = torch.randn_like(x)
    outs = []
    for v in ("0", "1"):
        monkeypatch.setenv("CNX_LNBWD_LINEARITY", v)
        blk.zero_grad(set_to_none=True)
        xi = x.clone().requires_grad_()
        blk(xi).backward(gy)
        outs.append((xi.grad, {n: p.grad.clone() for n, p in blk.named_parameters()}))
    (dx0, g0), (dx1, g1) = outs
    assert _rel(dx1, dx0.double()) < BF16_REL
    for n in g0:
        assert _rel(g1[n], g0[n].double()) < BF16_REL, n


@pytest.mark.parametrize("shape", [(2, 9, 11, 64), (1, 14, 14, 128), (1, 12, 12, 256), (2, 14, 14, 512),
                                   (1, 7, 7, 1024), (1, 24, 20, 192)])
def test_ln_backward_by_linearity_matches_row_kernel_and_fp64(cnx_lib, shape):
    """cnx_mlp_bwd_dz_stats + cnx_mlp_bwd_dconv (LayerNorm backward from row sums taken in the dz
    GEMM's epilogue, applied in the fc1 data-gradient GEMM's epilogue) against the fp64 formula and
    against the round-1 sequence (dxhat GEMM -> row kernel)."""
    from image_classification_b200 import ops
    B, H, W, C = shape
    M = B * H * W
    dev = _dev()
    torch.manual_seed(C + H)
    xhat = torch.randn(M, C, device=dev)
    xhat = _bf((xhat - xhat.mean(1, keepdim=True)) / xhat.std(1, unbiased=False, keepdim=True))
    w1 = torch.randn(4 * C, C, device=dev) * 0.08
    params = [torch.randn(C, 1, 7, 7, device=dev), torch.rand(C, device=dev) + 0.5,
              torch.randn(C, device=dev) * 0.1, w1, torch.randn(4 * C, device=dev) * 0.1,
              torch.randn(C, 4 * C, device=dev) * 0.05, torch.randn(C, device=dev) * 0.1,
              torch.rand(C, device=dev)]
    fw = ops.prep_block_weights(*params)
    w1f = fw.w1f.double()
    assert _rel(fw.rb[:, 0], w1f.sum(1)) < 1e-5 and torch.equal(fw.rb[:, 1], -fw.b1f)
    z, _h = ops.fc1_gelu_fwd(xhat.bfloat16(), fw.w1f, fw.b1f)
    dy = _bf(torch.randn(M, C, device=dev) * 0.1).bfloat16()
    rstd = torch.rand(M, device=dev) + 0.5
    dz, db1f, stats = ops.mlp_bwd_dz_stats(dy, fw.w2fT, z, fw.rb)
    dz_old, db1f_old = ops.mlp_bwd_dz(dy, fw.w2fT, z)
    assert torch.equal(dz, dz_old) and _rel(db1f, db1f_old) < 1e-5
    dxhat64 = dz.double() @ w1f                                  # what the GEMM accumulates
    s1, s2 = dxhat64.sum(1), (dxhat64 * xhat.double()).sum(1)
    scale = dxhat64.abs().sum(1).max()
    assert float((stats[:, 0].double() - s1).abs().max() / scale) < 1e-5
    # s2 is taken from the bf16-rounded z: a 2^-9 perturbation per term of a 4C-term sum
    assert float((stats[:, 1].double() - s2).abs().max() / (dxhat64.abs() * xhat.double().abs()).sum(1).max()) < 4e-3
    dconv = ops.mlp_bwd_dconv(dz, fw.w1fT, xhat.bfloat16(), stats, rstd)
    ref = rstd.double()[:, None] * (dxhat64 - s1[:, None] / C - xhat.double() * s2[:, None] / C)
    assert _rel(dconv, ref) < BF16_REL
    # the round-1 sequence on the same inputs: both within bf16 tolerance of fp64, and of each other
    dxhat = ops.mlp_bwd_dxhat(dz, fw.w1fT)
    x = _bf(torch.randn(B, H, W, C, device=dev)).bfloat16()
    dx_old, dwt_old, dbias_old, dycs_old = ops.dwconv_ln_bwd(
        dxhat.view(B, H, W, C), xhat.bfloat16().view(B, H, W, C), rstd, x, dy.view(B, H, W, C), fw.wt_dw)
    dx_new, dwt_new, dbias_new, dycs_new = ops.dwconv_bwd(dconv.view(B, H, W, C), x, dy.view(B, H, W, C), fw.wt_dw)
    assert _rel(dx_new, dx_old.double()) < BF16_REL
    assert _rel(dwt_new, dwt_old.double()) < BF16_REL and _rel(dbias_new, dbias_old.double()) < BF16_REL
    assert _rel(dycs_new, dycs_old.double()) < 1e-5
