"""Minimal launch sequence for `ncu --set full`: one ConvNeXt block forward+backward at the
stage-1 and stage-3 shapes of BASELINE config 2 (B=128, 224x224), after one warm-up pass."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200.block import ConvNeXtBlockB200
dev = torch.device("cuda:0")
torch.manual_seed(0)
for (C, H) in [(128, 56), (512, 14)]:
    blk = ConvNeXtBlockB200(C).to(dev)
    x = torch.randn(128, C, H, H, device=dev).bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_()
    gy = torch.randn_like(x)
    for it in range(2):            # it=0 warm-up (skipped by ncu -s), it=1 profiled
        y = blk(x); y.backward(gy)
    torch.cuda.synchronize()
print("ncu target done")
