"""ncu target: dwconv+LN forward and backward at one stage shape of BASELINE config 2
(3 launches each; profile the last).  python tools/ncu_conv_target.py [stage 0..3]"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import ops
s = int(sys.argv[1]) if len(sys.argv) > 1 else 2
dev = torch.device("cuda:0")
C = (128, 256, 512, 1024)[s]; H = 56 >> s; B = 128
torch.manual_seed(0)
x = torch.randn(B, H, H, C, device=dev).bfloat16()
wt = torch.randn(49, C, device=dev) * 0.1
bias = torch.randn(C, device=dev)
dy = torch.randn(B, H, H, C, device=dev).bfloat16()
dxhat = torch.randn(B, H, H, C, device=dev).bfloat16()
for _ in range(3):
    xhat, mean, rstd = ops.dwconv_ln_fwd(x, wt, bias)
    ops.dwconv_ln_bwd(dxhat, xhat, rstd, x, dy, wt)
torch.cuda.synchronize()
print("ncu conv target done", s)
