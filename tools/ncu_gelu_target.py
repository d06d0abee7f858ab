"""ncu target: the two GELU-carrying GEMMs (fc1+GELU forward, fc2 data-gradient x GELU') at one
stage shape of BASELINE config 2, 3 launches each (profile the last of each).
python tools/ncu_gelu_target.py [stage 0..3] [epi_warps]"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import ops, _lib
s = int(sys.argv[1]) if len(sys.argv) > 1 else 2
ew = int(sys.argv[2]) if len(sys.argv) > 2 else 0
_lib.load().cnx_debug_gemm_epilogue_warps(ew)
dev = torch.device("cuda:0")
C = (128, 256, 512, 1024)[s]
H = 56 >> s
M = 128 * H * H
torch.manual_seed(0)
xh = torch.randn(M, C, device=dev).bfloat16()
w1 = (torch.randn(4 * C, C, device=dev) * 0.05).bfloat16()
w2T = (torch.randn(4 * C, C, device=dev) * 0.05).bfloat16()
b1 = torch.randn(4 * C, device=dev)
dy = torch.randn(M, C, device=dev).bfloat16()
for _ in range(3):
    z, h = ops.fc1_gelu_fwd(xh, w1, b1)
    dz, _ = ops.mlp_bwd_dz(dy, w2T, z)
torch.cuda.synchronize()
print("ncu gelu target done", s, M, C)
