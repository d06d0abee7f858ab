"""Development: time the MLP GEMM entry points at the BASELINE config 2 stage shapes (CUDA events,
L2 flushed, median of 20).  python tools/gemm_quick.py [--stages 2] [--epi-warps 0] [--tag x]"""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import ops, _lib
from tools.microbench import timeit

ap = argparse.ArgumentParser()
ap.add_argument("--stages", default="0,1,2,3")
ap.add_argument("--epi-warps", type=int, default=0)
ap.add_argument("--tag", default="")
ap.add_argument("--ring", type=int, default=-1, help="operand ring depth (cnx_debug_gemm_tuning stages)")
ap.add_argument("--which", default="fc1,dz,fc2,dxhat,wgrad")
args = ap.parse_args()
_lib.load().cnx_debug_gemm_epilogue_warps(args.epi_warps)
_lib.load().cnx_debug_gemm_tuning(args.ring, -1, -1, -1)
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
which = args.which.split(",")
out = {}
for s in [int(v) for v in args.stages.split(",")]:
    C = (128, 256, 512, 1024)[s]; H = 56 >> s; M = 128 * H * H
    torch.manual_seed(s)
    xh = torch.randn(M, C, device=dev).bfloat16()
    x = torch.randn(M, C, device=dev).bfloat16()
    w1 = (torch.randn(4 * C, C, device=dev) * 0.05).bfloat16(); w1T = w1.t().contiguous()
    w2 = (torch.randn(C, 4 * C, device=dev) * 0.05).bfloat16(); w2T = w2.t().contiguous()
    b1 = torch.randn(4 * C, device=dev); b2 = torch.randn(C, device=dev)
    dy = torch.randn(M, C, device=dev).bfloat16()
    z, h = ops.fc1_gelu_fwd(xh, w1, b1)
    dz, _ = ops.mlp_bwd_dz(dy, w2T, z)
    fns = {"fc1": lambda: ops.fc1_gelu_fwd(xh, w1, b1), "dz": lambda: ops.mlp_bwd_dz(dy, w2T, z),
           "fc2": lambda: ops.fc2_residual_fwd(h, w2, b2, x), "dxhat": lambda: ops.mlp_bwd_dxhat(dz, w1T),
           "wgrad": lambda: ops.mlp_bwd_wgrad(dy, h, dz, xh)}
    for k in which:
        out[f"s{s + 1}_{k}"] = round(timeit(fns[k], iters=20, flush=flush) * 1e3, 1)
    del xh, x, z, h, dz, dy
    torch.cuda.empty_cache()
print(args.tag or os.environ.get("CNX_LIB_PATH", "current"), json.dumps(out), flush=True)
