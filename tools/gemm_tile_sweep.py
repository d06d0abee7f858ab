"""fc2 / dxhat / fc1 / dz at the stage-3 and stage-4 shapes of a given batch with forced column
tiles and pair modes: python tools/gemm_tile_sweep.py --batch 64"""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import _lib, ops
from tools.microbench import timeit

ap = argparse.ArgumentParser(); ap.add_argument("--batch", type=int, default=64); args = ap.parse_args()
dev = torch.device("cuda:0"); lib = _lib.load()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for (C, H) in ((512, 14), (1024, 7), (256, 28), (128, 56)):
    M = args.batch * H * H
    x = torch.randn(M, C, device=dev).bfloat16(); h = torch.randn(M, 4 * C, device=dev).bfloat16()
    w1 = (torch.randn(4 * C, C, device=dev) * 0.05).bfloat16(); w2 = (torch.randn(C, 4 * C, device=dev) * 0.05).bfloat16()
    b1 = torch.randn(4 * C, device=dev); b2 = torch.randn(C, device=dev)
    w1T = w1.t().contiguous(); w2T = w2.t().contiguous()
    fns = {"fc2_residual": lambda: ops.fc2_residual_fwd(h, w2, b2, x), "dxhat": lambda: ops.mlp_bwd_dxhat(h, w1T),
           "fc1_gelu": lambda: ops.fc1_gelu_fwd(x, w1, b1), "dz": lambda: ops.mlp_bwd_dz(x, w2T, h),
           "wgrad": lambda: ops.mlp_bwd_wgrad(x, h, h, x)}
    for name, fn in fns.items():
        row = {}
        for pair in (-1, 0, 1):
            for bn in (-1, 128, 256):
                lib.cnx_debug_gemm_tuning(-1, pair, bn, -1)
                row[f"p{pair}_bn{bn}"] = round(timeit(fn, flush=flush) * 1e3, 1)
        lib.cnx_debug_gemm_tuning(-1, -1, -1, -1)
        print(json.dumps({"C": C, "M": M, "kernel": name, **row}), flush=True)
