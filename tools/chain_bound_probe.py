"""Probe for the chained fc1 -> GELU -> fc2 kernel: what is the most such a
kernel could save at the stage-1 / stage-2 shapes of BASELINE config 2?

A chained kernel keeps the [M,4C] hidden tile on chip between the two GEMMs, but backward still
needs z (GELU') and h (dW2f = dy^T h), so it must WRITE both and read xhat and x once and write y:
bytes = (xhat + x + y) + (z + h).  This script times
  (a) the two shipped launches (cnx_fc1_gelu_fwd + cnx_fc2_residual_fwd),
  (b) a pure streaming pass that moves exactly the chained kernel's mandatory bytes (device
      copies of the same sizes: reads xhat, x; writes z, h, y) -- a floor no chained kernel beats,
  (c) the same for the two-launch form (adds the read of h by fc2).
python tools/chain_bound_probe.py  ->  one JSON line per stage
"""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import ops
from tools.microbench import timeit

dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for s, (C, H) in enumerate(((128, 56), (256, 28), (512, 14))):
    M = 128 * H * H
    xh = torch.randn(M, C, device=dev).bfloat16(); x = torch.randn(M, C, device=dev).bfloat16()
    w1 = (torch.randn(4 * C, C, device=dev) * 0.05).bfloat16(); w2 = (torch.randn(C, 4 * C, device=dev) * 0.05).bfloat16()
    b1 = torch.randn(4 * C, device=dev); b2 = torch.randn(C, device=dev)
    z, h = ops.fc1_gelu_fwd(xh, w1, b1)
    e = M * C * 2                                  # bytes of one [M, C] bf16 tensor
    a2 = torch.empty(2 * e, dtype=torch.uint8, device=dev); b2e = torch.empty(2 * e, dtype=torch.uint8, device=dev)
    w7 = torch.empty(7 * e, dtype=torch.uint8, device=dev)
    r4 = torch.empty(4 * e // 4, dtype=torch.float32, device=dev)

    def shipped():
        zz, hh = ops.fc1_gelu_fwd(xh, w1, b1)
        ops.fc2_residual_fwd(hh, w2, b2, x)

    def chained_floor():          # reads 2e (xhat, x), writes 9e (z, h: 4e each, y: e): contiguous streams
        a2.copy_(b2e)             # read 2e + write 2e
        w7.fill_(1)               # write 7e

    def two_launch_floor():       # + fc2 reads h (4e)
        chained_floor()
        r4.sum()

    r = {"stage": s + 1, "C": C, "M": M,
         "shipped_fc1_plus_fc2_us": round(timeit(shipped, iters=20, flush=flush) * 1e3, 1),
         "chained_mandatory_bytes_MB": round((3 * e + 8 * e) / 1e6, 1),
         "chained_floor_us": round(timeit(chained_floor, iters=20, flush=flush) * 1e3, 1),
         "two_launch_bytes_MB": round((3 * e + 12 * e) / 1e6, 1),
         "two_launch_floor_us": round(timeit(two_launch_floor, iters=20, flush=flush) * 1e3, 1)}
    r["max_saving_us"] = round(r["shipped_fc1_plus_fc2_us"] - r["chained_floor_us"], 1)
    print(json.dumps(r), flush=True)
