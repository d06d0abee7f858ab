"""One small launch of every kernel family, with a coarse check.  Meant to be run under
`compute-sanitizer --tool memcheck` BEFORE the plain test suite, so an out-of-bounds access is
reported by the tool instead of faulting the GPU."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import _lib, ops, mix, ema, loss
from image_classification_b200.block import ConvNeXtBlockB200

dev = torch.device("cuda:0")
rel = lambda a, b: float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))
def flag(tag):
    torch.cuda.synchronize()
    f = _lib.load().cnx_debug_timeout_flag()
    print(f"[{tag}] timeout_flag={f}", flush=True)
    return f

which = sys.argv[1] if len(sys.argv) > 1 else "all"
torch.manual_seed(0)
if which in ("all", "glue"):
    x = torch.randn(4, 3, 8, 8, device=dev); lab = torch.randint(0, 5, (4,), device=dev)
    out, t = mix.BatchMixUpCutMix(0.2, 1.0, 1.0, 5)(x, lab, rng=np.random.RandomState(0))
    l = loss.SoftTargetCrossEntropyB200(0.1)(torch.randn(4, 5, device=dev), t)
    m = torch.nn.Linear(9, 1000).to(dev); e = ema.EMAB200(m, 0.99); e.update()
    a = ops.nhwc_bf16_to_nchw_f32(ops.nchw_f32_to_nhwc_bf16(torch.randn(2, 40, 5, 7, device=dev)))
    flag("glue"); print("glue ok", float(l))
if which in ("all", "gemm"):
    for (M, N, K, amn, bmn) in [(128, 128, 64, 0, 0), (200, 136, 200, 0, 0), (136, 264, 200, 1, 1)]:
        A = torch.randn(M, K, device=dev).bfloat16(); B = torch.randn(N, K, device=dev).bfloat16()
        ref = A.double() @ B.double().t()
        D = ops.gemm_bf16(A.t().contiguous() if amn else A, B.t().contiguous() if bmn else B,
                          a_mn_major=bool(amn), b_mn_major=bool(bmn), epilogue=ops.EPI_STORE_F32)
        f = flag(f"gemm {M}x{N}x{K} mn={amn}{bmn}")
        print("   rel err", rel(D, ref), flush=True)
        if f: sys.exit(3)
if which in ("all", "dw"):
    for (B_, H, W, C) in [(1, 7, 7, 64), (1, 9, 11, 256)]:
        x = torch.randn(B_, H, W, C, device=dev).bfloat16()
        wt = torch.randn(49, C, device=dev) * 0.1; b = torch.randn(C, device=dev)
        xhat, mean, rstd = ops.dwconv_ln_fwd(x, wt, b)
        f = flag(f"dwfwd C={C}")
        y = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), wt.t().reshape(C, 1, 7, 7), b, padding=3, groups=C).permute(0, 2, 3, 1)
        ref = torch.nn.functional.layer_norm(y, (C,), eps=1e-6)
        print("   xhat rel err", rel(xhat, ref), flush=True)
        dx, dwt, db, dycs = ops.dwconv_ln_bwd(torch.randn_like(xhat), xhat, rstd, x, torch.randn_like(xhat), wt)
        f |= flag(f"dwbwd C={C}")
        print("   dx finite", bool(torch.isfinite(dx.float()).all()), flush=True)
        if f: sys.exit(3)
if which in ("all", "block"):
    blk = ConvNeXtBlockB200(128).to(dev)
    x = torch.randn(2, 128, 7, 7, device=dev).bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_()
    y = blk(x); y.float().sum().backward()
    flag("block"); print("block ok", bool(torch.isfinite(x.grad.float()).all()))
print("sanity done")
