"""Per-kernel timings at BASELINE config shapes (CUDA events, L2 flushed between iterations).

Development tool: prints one line per kernel with the achieved algorithmic GB/s or TFLOP/s
(SURVEY.md §8d accounting).  `python tools/microbench.py [--batch 128] [--res 224]`.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import ops  # noqa: E402


def timeit(fn, iters=10, warmup=3, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--res", type=int, default=224)
    ap.add_argument("--stages", type=str, default="0,1,2,3")
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--epi-warps", type=int, default=0, help="force 8 / 16 GEMM epilogue warps")
    ap.add_argument("--out", type=str, default="gpurun_out/microbench.json")
    args = ap.parse_args()
    from image_classification_b200 import _lib
    _lib.load().cnx_debug_gemm_epilogue_warps(args.epi_warps)
    dev = torch.device("cuda:0")
    dims = (128, 256, 512, 1024)
    flush = None if args.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []
    for s in [int(v) for v in args.stages.split(",")]:
        C = dims[s]
        H = W = args.res // 4 // (2 ** s)
        B = args.batch
        M = B * H * W
        E = M * C
        torch.manual_seed(s)
        x = torch.randn(B, H, W, C, device=dev).bfloat16()
        wt = torch.randn(49, C, device=dev) * 0.1
        bias = torch.randn(C, device=dev)
        w1 = (torch.randn(4 * C, C, device=dev) * 0.05).bfloat16()
        w1T = w1.t().contiguous()
        w2 = (torch.randn(C, 4 * C, device=dev) * 0.05).bfloat16()
        w2T = w2.t().contiguous()
        b1 = torch.randn(4 * C, device=dev)
        b2 = torch.randn(C, device=dev)
        xhat, mean, rstd = ops.dwconv_ln_fwd(x, wt, bias)
        x2d, xh2d = x.view(M, C), xhat.view(M, C)
        z, h = ops.fc1_gelu_fwd(xh2d, w1, b1)
        dy = torch.randn(M, C, device=dev).bfloat16()
        dz, _ = ops.mlp_bwd_dz(dy, w2T, z)
        dxhat = ops.mlp_bwd_dxhat(dz, w1T)
        gemm_flop = 8.0 * M * C * C   # one [M,C]x[C,4C] GEMM = 2*M*C*4C

        def rec(name, ms, gbytes=None, flop=None):
            r = dict(stage=s + 1, C=C, H=H, M=M, kernel=name, ms=round(ms, 4))
            if gbytes is not None:
                r["GB/s"] = round(gbytes / ms * 1e3, 1)
            if flop is not None:
                r["TFLOP/s"] = round(flop / ms * 1e-9, 1)
            rows.append(r)
            print(json.dumps(r), flush=True)

        t = timeit(lambda: ops.dwconv_ln_fwd(x, wt, bias), flush=flush)
        rec("dwconv_ln_fwd", t, gbytes=2 * E * 2 / 1e9)
        t = timeit(lambda: ops.dwconv_ln_bwd(dxhat.view(B, H, W, C), xhat, rstd, x,
                                             dy.view(B, H, W, C), wt), flush=flush)
        rec("dwconv_ln_bwd", t, gbytes=5 * E * 2 / 1e9)
        t = timeit(lambda: ops.fc1_gelu_fwd(xh2d, w1, b1), flush=flush)
        rec("fc1_gelu_fwd", t, flop=gemm_flop)
        t = timeit(lambda: ops.fc2_residual_fwd(h, w2, b2, x2d), flush=flush)
        rec("fc2_residual_fwd", t, flop=gemm_flop)
        t = timeit(lambda: ops.mlp_bwd_dz(dy, w2T, z), flush=flush)
        rec("mlp_bwd_dz", t, flop=gemm_flop)
        t = timeit(lambda: ops.mlp_bwd_dxhat(dz, w1T), flush=flush)
        rec("mlp_bwd_dxhat", t, flop=gemm_flop)
        t = timeit(lambda: ops.mlp_bwd_wgrad(dy, h, dz, xh2d), flush=flush)
        rec("mlp_bwd_wgrad(x2)", t, flop=2 * gemm_flop)
        # library yardsticks (NOT the product path): cuBLAS for the same GEMM shape
        t = timeit(lambda: torch.matmul(xh2d, w1.t()), flush=flush)
        rec("cublas_fc1_shape", t, flop=gemm_flop)
        t = timeit(lambda: torch.matmul(h, w2.t()), flush=flush)
        rec("cublas_fc2_shape", t, flop=gemm_flop)
        del x, xhat, z, h, dz, dxhat, dy
        torch.cuda.empty_cache()
    os.makedirs("gpurun_out", exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
