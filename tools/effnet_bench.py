"""EfficientNet-B2 path (BASELINE config 5) on one B200: per-kernel timings of the
BatchNorm+SiLU / squeeze-excite kernels at the network's shapes (CUDA events, L2 flushed between
iterations, algorithmic bytes per csrc/bnse.cu) and the training step of the older reference loop
(/root/reference/previous/V3-efb2/train_advanced_v3.0.py:72-153: autocast forward, hard-label CE,
scaled backward, unscale, clip 1.0, AdamW, scaler update) on torchvision's efficientnet_b2 with the
library modules swapped in (eager; with the fused optimizer step; with forward+backward replayed
as one CUDA graph), next to the same model on PyTorch's own bf16 channels_last kernels.

Development / measurement tool: `python tools/effnet_bench.py [--batch 48] [--res 260] [--steps 10]`.
Prints one JSON document.
"""
from __future__ import annotations

import argparse
import copy
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import _lib, ops  # noqa: E402
from image_classification_b200.efficientnet import swap_efficientnet  # noqa: E402


def timeit(fn, iters=10, warmup=3, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def kernel_rows(B, dev, flush, hbm_peak, once=False):
    """once: every kernel exactly once at two shapes (the ncu --set full target)."""
    rows = []
    shapes = [(96, 130), (144, 65), (288, 33), (528, 17), (720, 17), (1248, 9), (2112, 9)]
    for C, H in ([(96, 130), (528, 17)] if once else shapes):
        E = B * H * H * C
        x = torch.randn(B, H, H, C, device=dev).bfloat16()
        g = torch.randn(B, H, H, C, device=dev).bfloat16()
        w = torch.rand(C, device=dev) + 0.5
        b = torch.randn(C, device=dev) * 0.1
        gate = torch.rand(B, C, device=dev)
        dpool = torch.randn(B, C, device=dev)
        _, mean, rstd, _ = ops.bn_act_fwd(x, w, b, None, None, training=True, momentum=0.1,
                                          eps=1e-5, act=1)
        cases = {
            "bn_silu_fwd(+pool)": (3 * E * 2, lambda: ops.bn_act_fwd(
                x, w, b, None, None, training=True, momentum=0.1, eps=1e-5, act=1, want_pool=True)),
            "bn_silu_bwd": (5 * E * 2, lambda: ops.bn_act_bwd(g, x, w, b, mean, rstd, 1)),
            "se_scale_fwd": (2 * E * 2, lambda: ops.se_scale_fwd(x, gate)),
            "se_bwd_reduce": (2 * E * 2, lambda: ops.se_scale_bwd_reduce(g, x)),
            "se_bwd_apply": (2 * E * 2, lambda: ops.se_scale_bwd_apply(g, gate, dpool)),
        }
        for name, (nbytes, fn) in cases.items():
            if once:
                fn()
                torch.cuda.synchronize()
                continue
            ms = timeit(fn, flush=flush)
            gbs = nbytes / ms / 1e6
            rows.append({"kernel": name, "C": C, "H": H, "MB": round(nbytes / 1e6, 1),
                         "us": round(ms * 1e3, 1), "GBps": round(gbs, 1),
                         "frac_hbm": round(gbs / hbm_peak, 3)})
    return rows


def train_step_ms(model, x, labels, steps, warmup, fused=False, graphed=False, world=1):
    """The reference V3 step (train_advanced_v3.0.py:96-113) with bf16 autocast.  fused: the same
    sequence through train_step.TrainStep (accumulation 1, no EMA): library soft-CE kernel and
    unscale + clip + AdamW as two launches (optim_step.FusedOptimStep)."""
    opt = torch.optim.AdamW(model.parameters(), lr=2e-4, weight_decay=1e-2)     # config.py:17-18
    crit = torch.nn.CrossEntropyLoss()                                          # :449
    scaler = torch.amp.GradScaler("cuda")
    if fused:
        from image_classification_b200.loss import SoftTargetCrossEntropyB200
        from image_classification_b200.train_step import StepConfig, TrainStep
        ts = TrainStep(model, SoftTargetCrossEntropyB200(0.0), opt, None, scaler, None,
                       StepConfig(gradient_accumulation_steps=1, gradient_clip_val=1.0))

    if graphed:
        from image_classification_b200.graphed import GraphedForwardBackward
        from image_classification_b200.loss import SoftTargetCrossEntropyB200
        from image_classification_b200.optim_step import FusedOptimStep
        reducer = None
        if world > 1:        # data parallel: bucketed NCCL all-reduce captured inside the graph
            from image_classification_b200.ddp import BucketedGradReducer
            reducer = BucketedGradReducer(list(model.parameters()), bucket_bytes=8 << 20)
        gfb = GraphedForwardBackward(model, SoftTargetCrossEntropyB200(0.0), scaler, x, labels,
                                     grad_reducer=reducer if os.environ.get("EFFNET_NCCL_OUTSIDE") is None else None)
        fo = FusedOptimStep(model, opt, scaler, None, 1.0)

    def step():
        if graphed:                      # one graph launch + the two optimizer launches
            loss = gfb(x, labels)
            if reducer is not None and gfb.reducer is None:     # A/B: all-reduce after the replay
                reducer.set_sync(True)
                reducer.finish()         # hooks do not fire in a replay: finish() sends every bucket
            fo.step()
            return loss
        if fused:
            return ts.micro_step(x, labels, 0)
        opt.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = model(x)
            loss = crit(out.float(), labels)
        scaler.scale(loss).backward()
        scaler.unscale_(opt)
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)                 # config.py:22
        scaler.step(opt)
        scaler.update()
        return loss

    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    e2e = None
    if graphed and world == 1:
        # end to end: every step's batch comes from pinned host memory (H2D on the copy stream, one
        # batch ahead) and every step's loss travels back to the host, inside the timed region
        from image_classification_b200.prefetch import AsyncScalarReader, DevicePrefetcher
        hx = [torch.randn(x.shape).contiguous(memory_format=torch.channels_last).pin_memory()
              for _ in range(2)]
        hy = [torch.randint(0, 44, labels.shape).pin_memory() for _ in range(2)]
        reader, losses = AsyncScalarReader(x.device), []
        torch.cuda.synchronize()
        s0, e0 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        nwarm = max(warmup, 3)       # the copy stream's staging blocks are allocated in these
        feed = ((hx[i % 2], hy[i % 2]) for i in range(nwarm + steps))
        for i, (xb, yb) in enumerate(DevicePrefetcher(feed, x.device)):
            if i == nwarm:
                s0.record()
            l = gfb(xb, yb)
            fo.step()
            if i >= nwarm:
                losses += reader.push(l)
        losses += reader.drain()
        e0.record()
        torch.cuda.synchronize()
        e2e = {"ms_per_step": round(s0.elapsed_time(e0) / steps, 3),
               "h2d_bytes_per_step": hx[0].numel() * 4 + hy[0].numel() * 8, "d2h_bytes_per_step": 4,
               "losses_read": len(losses)}
    if world > 1:
        torch.distributed.barrier()
        torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps):
        loss = step()
    e.record()
    torch.cuda.synchronize()
    ms = torch.tensor([s.elapsed_time(e) / steps], device=x.device)
    if world > 1:
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)     # slowest rank
    return float(ms), float(loss.detach()), e2e


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=48)          # previous/V3-efb2/config.py:14
    ap.add_argument("--res", type=int, default=260)           # :12
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--skip-kernels", action="store_true")
    ap.add_argument("--only-kernels", action="store_true", help="per-kernel table only")
    ap.add_argument("--ncu", action="store_true", help="each kernel once at two shapes (ncu target)")
    ap.add_argument("--profile", action="store_true",
                    help="torch.profiler kernel table of 3 eager steps of the library arm (diagnosis)")
    args = ap.parse_args()
    import torchvision
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:            # torchrun: one process per GPU, NCCL over NVLink
        torch.distributed.init_process_group("nccl", device_id=dev)
        args.skip_kernels = True
    peaks = {}
    pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    hbm = float(peaks.get("hbm_gbs", 6500.0))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = {"batch": args.batch, "res": args.res, "hbm_peak_gbs": hbm,
           "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback 6500"}
    if args.ncu:
        kernel_rows(args.batch, dev, None, hbm, once=True)
        return
    if not args.skip_kernels:
        out["kernels"] = kernel_rows(args.batch, dev, flush, hbm)
    if args.only_kernels:
        print(json.dumps(out))
        return
    torch.manual_seed(0)
    base = torchvision.models.efficientnet_b2(weights=None, num_classes=44).to(dev).train()
    for m in base.modules():                  # timm's create_model defaults: no drop-path
        if isinstance(m, torchvision.ops.StochasticDepth):
            m.p = 0.0
    torch.manual_seed(1 + rank)               # same weights on every rank, different batches
    x = torch.randn(args.batch, 3, args.res, args.res, device=dev).contiguous(
        memory_format=torch.channels_last)
    labels = torch.randint(0, 44, (args.batch,), device=dev)
    out["n_gpus"] = world
    if args.profile:
        from torch.profiler import ProfilerActivity, profile
        model = copy.deepcopy(base)
        swap_efficientnet(model)
        model = model.to(memory_format=torch.channels_last)
        train_step_ms(model, x, labels, 2, 2, fused=True)
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            train_step_ms(model, x, labels, 3, 0, fused=True)      # (ms, loss, e2e) unused
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45,
                                        max_name_column_width=70))
        return
    arms = ("b200_graphed",) if world > 1 else ("b200", "b200_fused_optim", "b200_graphed",
                                                "torch_bf16_channels_last")
    for arm in arms:
        model = copy.deepcopy(base)
        if arm.startswith("b200"):
            out["swapped"] = swap_efficientnet(model)
        model = model.to(memory_format=torch.channels_last)
        n0 = _lib.launch_count
        ms, loss, e2e = train_step_ms(model, x, labels, args.steps, args.warmup,
                                 fused=arm == "b200_fused_optim", graphed=arm == "b200_graphed",
                                 world=world)
        out[arm] = {"ms_per_step": round(ms, 3),
                    "images_per_s": round(world * args.batch / ms * 1e3, 1),
                    "loss": round(loss, 4),
                    "lib_launches_per_step": (_lib.launch_count - n0) // (
                        args.steps * (2 if e2e is not None else 1) + args.warmup * (2 if e2e is not None else 1))}
        if e2e is not None:
            e2e["images_per_s"] = round(args.batch / e2e["ms_per_step"] * 1e3, 1)
            out[arm]["e2e"] = e2e
        del model
        torch.cuda.empty_cache()
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
