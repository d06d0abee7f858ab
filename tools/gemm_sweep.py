"""Same-process A/B sweeps of the GEMM's tiling knobs (cnx_debug_gemm_tuning): ring depth, CTA
pairs, column-tile width, split-K.  `python tools/gemm_sweep.py [stage]` prints one JSON line per
shape with microseconds per configuration."""
import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import ops, _lib
from tools.microbench import timeit

dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def tune(stages=-1, pair=-1, bn=-1, splits=-1):
    _lib.load().cnx_debug_gemm_tuning(stages, pair, bn, splits)


def us(fn):
    return round(timeit(fn, iters=12, flush=flush) * 1e3, 1)


stage = int(sys.argv[1]) if len(sys.argv) > 1 else 3
C, H = [(128, 56), (256, 28), (512, 14), (1024, 7)][stage - 1]
M = 128 * H * H
x = torch.randn(M, C, device=dev).bfloat16()
hh = torch.randn(M, 4 * C, device=dev).bfloat16()
w1 = (torch.randn(4 * C, C, device=dev) * 0.05).bfloat16()      # fc1: N=4C, K=C
w2 = (torch.randn(C, 4 * C, device=dev) * 0.05).bfloat16()      # fc2: N=C, K=4C
b1 = torch.randn(4 * C, device=dev)
b2 = torch.randn(C, device=dev)
z = torch.randn(M, 4 * C, device=dev).bfloat16()

shapes = {
    "plain_N4C_K1C": lambda: ops.gemm_bf16(x, w1),
    "plain_N1C_K4C": lambda: ops.gemm_bf16(hh, w2),
    "gelu_N4C_K1C": lambda: ops.gemm_bf16(x, w1, epilogue=ops.EPI_BIAS_GELU, bias=b1),
    "gelugrad_N4C_K1C": lambda: ops.gemm_bf16(x, w1, epilogue=ops.EPI_GELU_GRAD, R=z),
    "resid_N1C_K4C": lambda: ops.gemm_bf16(hh, w2, epilogue=ops.EPI_BIAS_RESIDUAL, bias=b2, R=x),
}
for name, fn in shapes.items():
    r = {"stage": stage, "shape": name}
    for pair in (0, 1):
        for bn in (256, 128):
            for st in (2, 3, 4, 5, 6, 8):
                tune(stages=st, pair=pair, bn=bn)
                r[f"p{pair}_bn{bn}_s{st}"] = us(fn)
    tune()
    r["auto"] = us(fn)
    print(json.dumps(r), flush=True)

# weight gradient: D[4C, C] += hh^T x  (both operands MN-major, K = tokens)
r = {"stage": stage, "shape": "wgrad_M4C_N1C"}
fn = lambda: ops.gemm_bf16(hh, x, a_mn_major=True, b_mn_major=True, epilogue=ops.EPI_STORE_F32)
base = us(lambda: torch.zeros(4 * C, C, device=dev))
r["zeros_only"] = base
for pair in (0, 1):
    for st in (3, 4, 5, 7):
        for sp in (-1, 4, 5, 9, 18):
            tune(stages=st, pair=pair, splits=sp)
            r[f"p{pair}_s{st}_k{sp}"] = us(fn)
tune()
r["auto"] = us(fn)
r["cublas"] = us(lambda: torch.matmul(hh.t(), x))
print(json.dumps(r), flush=True)
