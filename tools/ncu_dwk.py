"""ncu target: the depthwise k x k kernels once at two EfficientNet-B2 shapes."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import ops
dev = torch.device("cuda:0")
for k, s, C, H in [(5, 1, 720, 17), (3, 1, 144, 65)]:
    x = torch.randn(48, H, H, C, device=dev).bfloat16()
    wt = torch.randn(k * k, C, device=dev) * 0.2
    y = ops.dwconv_kxk_fwd(x, wt, k, s)
    g = torch.randn_like(y)
    ops.dwconv_kxk_bwd(g, x, wt, k, s)
    torch.cuda.synchronize()
print("ok")
