"""Development: time cnx_dwconv_ln_fwd / cnx_dwconv_ln_bwd at the BASELINE config 2 stage shapes
(CUDA events, L2 flushed, median of 20).  python tools/conv_quick.py [--tag x] [--batch 128]"""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import ops
from tools.microbench import timeit
ap = argparse.ArgumentParser(); ap.add_argument("--tag", default=""); ap.add_argument("--batch", type=int, default=128)
args = ap.parse_args()
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
out = {}
for s, C in enumerate((128, 256, 512, 1024)):
    H = 56 >> s; B = args.batch
    torch.manual_seed(s)
    x = torch.randn(B, H, H, C, device=dev).bfloat16()
    wt = torch.randn(49, C, device=dev) * 0.1; bias = torch.randn(C, device=dev)
    dy = torch.randn(B, H, H, C, device=dev).bfloat16(); dxhat = torch.randn(B, H, H, C, device=dev).bfloat16()
    xhat, mean, rstd = ops.dwconv_ln_fwd(x, wt, bias)
    out[f"s{s+1}_fwd"] = round(timeit(lambda: ops.dwconv_ln_fwd(x, wt, bias), iters=20, flush=flush) * 1e3, 1)
    out[f"s{s+1}_bwd"] = round(timeit(lambda: ops.dwconv_ln_bwd(dxhat, xhat, rstd, x, dy, wt), iters=20, flush=flush) * 1e3, 1)
print(args.tag, json.dumps(out), flush=True)
