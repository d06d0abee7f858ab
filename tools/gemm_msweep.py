"""Fixed cost per GEMM launch: time the plain stage-3 GEMMs at 1/4 .. 4x the token count, with the
L2 flushed between iterations and with it left warm.  `python tools/gemm_msweep.py`"""
import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import ops, _lib
from tools.microbench import timeit

dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
C = 512
w1 = (torch.randn(4 * C, C, device=dev) * 0.05).bfloat16()
w2 = (torch.randn(C, 4 * C, device=dev) * 0.05).bfloat16()
for pair in (1, 0):
    _lib.load().cnx_debug_gemm_tuning(-1, pair, 256, -1)
    for name, K, w in (("N4C_K1C", C, w1), ("N1C_K4C", 4 * C, w2)):
        r = {"shape": name, "pair": pair}
        for f in (0.125, 0.25, 0.5, 1, 2, 4):
            M = int(25088 * f)
            a = torch.randn(M, K, device=dev).bfloat16()
            fn = lambda: ops.gemm_bf16(a, w)
            r[f"M{M}_cold"] = round(timeit(fn, iters=12, flush=flush) * 1e3, 1)
            r[f"M{M}_warm"] = round(timeit(fn, iters=12, flush=None) * 1e3, 1)
            del a
        print(json.dumps(r), flush=True)
