"""Probe: ConvNeXt-Base forward + soft-CE + backward of one 128-image micro-batch, eager launches
against one CUDA-graph replay (graphed.GraphedForwardBackward).  Answers whether the flagship step
(GPU-bound: the host enqueues a step in ~25 ms, the GPU needs ~44 ms) would gain from replaying.
`python tools/graph_probe.py [--batch 128] [--steps 10]`"""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import loss as loss_mod
from image_classification_b200.convnext import create_convnext, DeepSupervisionModelB200
from image_classification_b200.graphed import GraphedForwardBackward

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--steps", type=int, default=10)
args = ap.parse_args()
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = DeepSupervisionModelB200(create_convnext("convnext_base", 44), 44).to(dev).train()
for m in model.modules():
    if isinstance(m, torch.nn.Conv2d) and m.groups == 1:
        m.to(memory_format=torch.channels_last)
crit = loss_mod.SoftTargetCrossEntropyB200(0.1)
scaler = torch.amp.GradScaler("cuda")
x = torch.randn(args.batch, 3, 224, 224, device=dev)
t = torch.softmax(torch.randn(args.batch, 44, device=dev), dim=1)


def eager():
    for p in model.parameters():
        p.grad = None
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out = model(x)
        out = out[0] if isinstance(out, (list, tuple)) else out
        l = crit(out, t)
    scaler.scale(l).backward()
    return l


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(args.steps):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / args.steps


ms_eager = timed(eager)
gfb = GraphedForwardBackward(model, crit, scaler, x, t)
ms_graph = timed(lambda: gfb(x, t))
print(json.dumps({"batch": args.batch, "eager_ms": round(ms_eager, 3), "graphed_ms": round(ms_graph, 3),
                  "ratio": round(ms_eager / ms_graph, 4)}))
