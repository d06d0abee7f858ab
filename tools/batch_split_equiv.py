"""One GPU: gradients of a 16-image batch in ONE pass vs the mean of two 8-image passes (what two
data-parallel ranks compute), ConvNeXt-Base, against each other and against the fp32 oracle model.
Diagnosis tool for tools/ddp_equiv.py: `python tools/batch_split_equiv.py [--res 96]`."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("CNX_STRICT", "1")
import torch

from image_classification_b200.convnext import create_convnext
from image_classification_b200.loss import SoftTargetCrossEntropyB200
from oracle.model_oracle import build_oracle_model
from oracle.block_oracle import soft_ce_ref

ap = argparse.ArgumentParser()
ap.add_argument("--res", type=int, default=96)
ap.add_argument("--batch", type=int, default=8)
ap.add_argument("--oracle", action="store_true")
args = ap.parse_args()
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = create_convnext("convnext_base", 44).to(dev).train()
with torch.no_grad():
    for n, p in m.named_parameters():
        if n.endswith("gamma"):
            p.copy_(torch.rand_like(p) * 0.9 + 0.1)
b, R = args.batch, args.res
g = torch.Generator().manual_seed(42)
X = torch.randn(2 * b, 3, R, R, generator=g).to(dev)
T = torch.softmax(torch.randn(2 * b, 44, generator=g) * 2, 1).to(dev)
crit = SoftTargetCrossEntropyB200(0.1)


def grads(parts, scale):
    m.zero_grad(set_to_none=True)
    for x, t in parts:
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = crit(m(x), t) * scale
        loss.backward()
    return {n: p.grad.detach().clone() for n, p in m.named_parameters()}


g_full = grads([(X, T)], 1.0)
g_split = grads([(X[:b], T[:b]), (X[b:], T[b:])], 0.5)
g_full2 = grads([(X, T)], 1.0)


def cmp(a, c):
    worst, name = 0.0, ""
    for n in a:
        s = float(a[n].abs().max().clamp_min(1e-30))
        e = float((a[n] - c[n]).abs().max()) / s
        if e > worst:
            worst, name = e, n
    return worst, name


out = {"full_vs_split": cmp(g_full, g_split), "full_vs_full_again": cmp(g_full, g_full2)}
if args.oracle:
    torch.backends.cudnn.allow_tf32 = False
    o = build_oracle_model("convnext_base", 44).to(dev).train()
    o.load_state_dict(m.state_dict())
    soft_ce_ref(o(X), T, 0.1).backward()
    g_o = {n: p.grad for n, p in o.named_parameters()}
    out["full_vs_oracle"] = cmp(g_o, g_full)
    out["split_vs_oracle"] = cmp(g_o, g_split)
# per-stage first-block fc1 weight, to see where a difference starts
for n in g_full:
    if n.endswith("blocks.0.mlp.fc1.weight") or n.startswith("head.fc") or n.startswith("stem.0.w"):
        s = float(g_full[n].abs().max())
        out[n] = float((g_full[n] - g_split[n]).abs().max()) / s
print(json.dumps(out, indent=1))
