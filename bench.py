"""bench.py — ConvNeXt-Base 224x224 bf16 training throughput (images/s) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

One "step" = one optimizer step of the reference loop (/root/reference/train_advanced_v4.py:
215-237): 2 accumulation micro-batches of 128 images per GPU, each = device MixUp/CutMix ->
ConvNeXt-Base forward (36 libcnx_b200 blocks) -> label-smoothed soft-target CE -> backward; then
unscale / clip 1.0 / AdamW / scheduler / EMA.  `value` = images all ranks processed / max-over-
ranks device time with the raw batch already in HBM; `e2e` = same loop fed from pinned host
buffers through the public API (`prefetch.DevicePrefetcher`: H2D copy of every micro-batch on a
second stream, one batch ahead; D2H read of every step's loss through pinned memory — all inside
the timed region).  The headline is WEAK scaling on BASELINE config 2's per-GPU shape (global
batch = 256 * N) so rounds stay comparable; after the headline region the same process times the
other BASELINE configs (~10 steps each) and prints them under `configs`:
  cfg3 = global batch 1024 split over N ranks, accumulation 2 (/root/reference/config.py:37): micro
         256 / 128 / 64 at N = 2 / 4 / 8 — STRONG scaling; at N = 1 the 8-GPU rank's share (micro 64)
         is timed so the driver's 1-GPU record carries that shape too;
  cfg4 = 384x384, micro-batch 64 per GPU;
  cfg5 = EfficientNet-B2 260x260 batch 48 per GPU through swap_efficientnet +
         GraphedForwardBackward (/root/reference/previous/V3-efb2/config.py:10-15).

`--impl reference` times the reference's CPU PyTorch path for the same metric (oracle port of the
model, the reference loop order, bounded micro-batches) on the host cores of rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch

METRIC = "ConvNeXt-Base 224 bf16 train images/sec"
SM_COUNT = 148
UNIT = "images/s"
NUM_CLASSES = 44
MICRO_BATCH = 128
ACCUM = 2
RES = 224
DIMS = (128, 256, 512, 1024)
DEPTHS = (3, 3, 27, 3)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", type=str, default="b200", choices=["b200", "reference"])
    ap.add_argument("--res", type=int, default=RES)
    ap.add_argument("--micro-batch", type=int, default=MICRO_BATCH)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the cfg3 / cfg4 / cfg5 legs")
    ap.add_argument("--config-steps", type=int, default=10)
    return ap.parse_args()


def peaks():
    """Measured roofline denominators (driver-written) or the profiling guide's fallback."""
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm_gbs=d["hbm_gbs"], bf16_tflops=d["bf16_tflops"],
                    bf16_tflops_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_tflops_sustained=1400.0,
                source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 100 ms.  The process is started during
    warm-up (nvidia-smi needs ~0.2 s to come up) and `mark()` drops what it printed before the
    timed region, so every remaining sample was taken under the timed load."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def mark(self):
        self.before, self.lines = self.lines[-1:], []

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        # a timed region shorter than the sampling period: the last warm-up sample (same load)
        for ln in (self.lines or getattr(self, "before", [])):
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); smax.append(float(parts[1])); power.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# ---------------------------------------------------------------------------------------------
# CPU arm: the reference's PyTorch path on the host cores (oracle port; bounded micro-batches)
# ---------------------------------------------------------------------------------------------
def run_cpu_reference(steps: int, warmup: int, micro_batch: int, res: int, model_name: str):
    """-> (images/s, ms per step, cores, sample description)."""
    from oracle import loop_oracle   # the checker, timed here only as the stated CPU baseline
    from image_classification_b200.train_step import get_cosine_schedule_with_warmup
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model = loop_oracle.build_oracle_model(model_name, NUM_CLASSES, deep_supervision=True).train()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-4, weight_decay=1e-2)   # config.py:35-36
    sched = get_cosine_schedule_with_warmup(opt, 10, 1000, 1e-6)
    ema = loop_oracle.EMARef(model, 0.9997)
    g = torch.Generator().manual_seed(0)
    images = torch.randn(ACCUM, micro_batch, 3, res, res, generator=g)
    labels = torch.randint(0, NUM_CLASSES, (ACCUM, micro_batch), generator=g)
    rng = np.random.RandomState(0)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        loop_oracle.reference_cpu_step(model, opt, sched, ema, images, labels, accum=ACCUM,
                                       clip=1.0, label_smoothing=0.1, num_classes=NUM_CLASSES,
                                       rng=rng)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    total = sum(times)
    ips = steps * ACCUM * micro_batch / total
    sample = (f"{model_name} fp32 CPU, {steps} optimizer steps of {ACCUM}x{micro_batch} images at "
              f"{res}x{res} (reference loop order: per-sample MixUp/CutMix, CE(ls=0.1), "
              f"accumulate {ACCUM}, clip 1.0, AdamW, Python EMA), {warmup} warm-up")
    return ips, 1e3 * total / steps, cores, sample


def run_cpu_config1(warmup: int = 2, timed: int = 5):
    """BASELINE config 1 exactly: ConvNeXt-Tiny, 44 classes, 224x224, batch 8, fp32,
    random init, one forward + backward with CrossEntropyLoss(label_smoothing=0.1) on soft [8,44]
    targets (train_advanced_v4.py:626, config.py:31); 2 warm-up + 5 timed iterations, median."""
    from oracle import loop_oracle
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model = loop_oracle.build_oracle_model("convnext_tiny", NUM_CLASSES).train()
    crit = torch.nn.CrossEntropyLoss(label_smoothing=0.1)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(8, 3, 224, 224, generator=g)
    t = torch.softmax(torch.randn(8, NUM_CLASSES, generator=g) * 2, 1)
    ts = []
    for i in range(warmup + timed):
        t0 = time.perf_counter()
        model.zero_grad(set_to_none=True)
        crit(model(x), t).backward()
        if i >= warmup:
            ts.append(time.perf_counter() - t0)
    med = statistics.median(ts)
    return {"value": round(8 / med, 3), "unit": UNIT, "cores": cores, "kind": "port",
            "ms_per_iter": round(1e3 * med, 1),
            "sample": f"BASELINE config 1: convnext_tiny fp32 CPU, batch 8 at 224x224, forward + "
                      f"backward + CE(ls=0.1) on soft targets, {warmup} warm-up + {timed} timed, median"}


def main_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return 0
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    # bounded sample: a ConvNeXt-Base step of 2x8 images is ~seconds of CPU work; shrink the
    # micro-batch when many steps are requested so the whole run stays within a few minutes
    total = steps + warmup
    micro = 8 if total <= 30 else (4 if total <= 80 else 2)
    ips, ms, cores, sample = run_cpu_reference(steps, warmup, micro, args.res, "convnext_base")
    line = {
        "impl": "reference", "metric": METRIC, "value": round(ips, 3), "unit": UNIT,
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": round(ms, 2),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": f"ConvNeXt-Base 44-class {args.res}x{args.res} training step, "
                               f"reference CPU PyTorch path (bounded sample: micro-batch {micro})",
                   "micro_batch": micro, "accumulation": ACCUM, "optimizer": "AdamW", "ema": 0.9997},
        "cpu_baseline": {"value": round(ips, 3), "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": sample},
        "e2e": {"value": round(ips, 3), "unit": UNIT, "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------
def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant GEMM family, from
    the committed summary of an `ncu --set full` capture of the SHIPPED build
    (profiles/r02_ncu_gemm_traffic.json, written by tools/ncu_traffic.py with the git hash of the
    build it profiled).  None when no capture is committed."""
    path = os.path.join(ROOT, "profiles", "r02_ncu_gemm_traffic.json")
    if not os.path.exists(path):
        return None, None
    d = json.load(open(path))
    return d.get("bytes_per_launch_mean"), (f"{d.get('what', '')} (profiles/r02_ncu_gemm_traffic.json, "
                                            f"build {d.get('git', '?')}); algorithmic bytes of the same "
                                            f"launches {d.get('algorithmic_bytes_mean')}")


class B200Trainer:
    """Everything one rank needs to run optimizer steps through the repo's public API."""

    def __init__(self, device, world, rank, micro_batch, res):
        from image_classification_b200 import ddp, ema, loss, mix, train_step
        from image_classification_b200.convnext import create_convnext, DeepSupervisionModelB200
        self.device, self.world, self.rank = device, world, rank
        self.micro_batch, self.res = micro_batch, res
        torch.manual_seed(0)                      # identical replicas on every rank
        # create_model(): timm convnext_base wrapped for deep supervision (config.py:22, :494-495);
        # the loop keeps outputs[0], the three auxiliary heads run forward only
        self.model = DeepSupervisionModelB200(create_convnext("convnext_base", NUM_CLASSES),
                                              NUM_CLASSES).to(device).train()
        if os.environ.get("CNX_NO_AUX"):          # A/B switch: skip the dead auxiliary heads (f-1)
            self.model.compute_aux = False
        for m in self.model.modules():            # stem / downsample convs produce NHWC directly
            if isinstance(m, torch.nn.Conv2d) and m.groups == 1:
                m.to(memory_format=torch.channels_last)
        self.params = [p for p in self.model.parameters() if p.requires_grad]
        self.opt = torch.optim.AdamW(self.params, lr=1e-4, weight_decay=1e-2, fused=True)
        self.sched = train_step.get_cosine_schedule_with_warmup(self.opt, 10, 100000, 1e-6)
        self.scaler = torch.amp.GradScaler("cuda")            # kept: reference control flow (:432)
        self.ema = ema.EMAB200(self.model, 0.9997)            # config.py:47
        self.criterion = loss.SoftTargetCrossEntropyB200(0.1)  # config.py:31
        self.mixer = mix.BatchMixUpCutMix(0.2, 1.0, 0.5, NUM_CLASSES)   # config.py:16-17, :593-599
        self.rng = np.random.RandomState(1234 + rank)
        # the reducer finds the never-used auxiliary heads on its first synchronised backward
        bucket_mb = int(os.environ.get("CNX_BUCKET_MB", "32"))       # A/B switch
        self.reducer = (ddp.BucketedGradReducer(self.params, bucket_bytes=bucket_mb << 20)
                        if world > 1 else None)
        cfg = train_step.StepConfig(gradient_accumulation_steps=ACCUM, gradient_clip_val=1.0,
                                    num_classes=NUM_CLASSES)
        # CNX_NO_FUSED_OPTIM=1: torch's unscale_/clip/AdamW/update + EMAB200.update (A/B switch)
        fused = False if os.environ.get("CNX_NO_FUSED_OPTIM") else None
        self.step_obj = train_step.TrainStep(self.model, self.criterion, self.opt, self.sched,
                                             self.scaler, self.ema, cfg, self.reducer,
                                             use_fused_optim=fused)
        # forward + CE + backward of the two micro-batch positions as two CUDA graphs
        # (graphed.GraphedTrainStep); CNX_NO_GRAPH=1: every launch enqueued from Python (A/B switch)
        self.graphed = None
        if not os.environ.get("CNX_NO_GRAPH"):
            from image_classification_b200.graphed import GraphedTrainStep
            self.graphed = GraphedTrainStep(self.step_obj, (micro_batch, 3, res, res),
                                            (micro_batch, NUM_CLASSES), device)
        g = torch.Generator().manual_seed(100 + rank)
        # synthetic "normalised pixels" (dataset.py:233-236) and hard labels, pinned on the host
        self.host_images = [torch.randn(micro_batch, 3, res, res, generator=g).pin_memory()
                            for _ in range(ACCUM)]
        self.host_labels = [torch.randint(0, NUM_CLASSES, (micro_batch,), generator=g).pin_memory()
                            for _ in range(ACCUM)]
        self.dev_images = [t.to(device) for t in self.host_images]
        self.dev_labels = [t.to(device) for t in self.host_labels]
        self.batch_idx = 0

    def _micro(self, images, labels, eager=False):
        if self.graphed is not None and not eager:   # the mix kernel writes the graphs' static inputs
            mixed, targets = self.mixer(images, labels, rng=self.rng, out=self.graphed.images,
                                        targets_out=self.graphed.targets)
            loss = self.graphed.micro_step(mixed, targets, self.batch_idx)
        else:
            mixed, targets = self.mixer(images, labels, rng=self.rng)
            loss = self.step_obj.micro_step(mixed, targets, self.batch_idx)
        self.batch_idx += 1
        return loss

    def step_resident(self, eager=False):
        """inputs already in HBM; eager=True: every launch enqueued from Python (the per-kernel
        event instrumentation of profile_kernels needs the individual library calls)"""
        for i in range(ACCUM):
            loss = self._micro(self.dev_images[i], self.dev_labels[i], eager)
        return loss

    def run_e2e(self, steps: int) -> list[float]:
        """`steps` optimizer steps fed from pinned host memory through the package's input
        pipeline: the H2D copy of every micro-batch (second stream, one batch ahead) and a D2H read
        of every step's loss (pinned slot, read one step later) are all inside the caller's timed
        region; returns the losses."""
        from image_classification_b200.prefetch import AsyncScalarReader, DevicePrefetcher

        def host_batches():
            for _ in range(steps):
                for i in range(ACCUM):
                    yield self.host_images[i], self.host_labels[i]

        reader = AsyncScalarReader(self.device)
        losses = []
        for j, (images, labels) in enumerate(DevicePrefetcher(host_batches(), self.device)):
            loss = self._micro(images, labels)
            if (j + 1) % ACCUM == 0:
                losses += reader.push(loss)
        losses += reader.drain()
        return losses

    @property
    def h2d_bytes_per_step(self):
        return ACCUM * (self.host_images[0].numel() * 4 + self.host_labels[0].numel() * 8)

    def replica_checksums(self):
        """(sum of parameters, sum of |parameters|, sum of EMA shadows) in float64."""
        with torch.no_grad():
            ps = torch.stack([p.detach().double().sum() for p in self.params]).sum()
            pa = torch.stack([p.detach().double().abs().sum() for p in self.params]).sum()
            es = torch.stack([s.double().sum() for s in self.ema.shadow.values()]).sum()
        return torch.stack([ps, pa, es])

    def close(self):
        self.step_obj.close()
        for h in (self.reducer._hooks if self.reducer is not None else []):
            h.remove()


def timed_region(fn, steps, world, device):
    """barrier + sync, K steps between CUDA events on the current stream, max over ranks (ms)."""
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps):
        fn()
    e.record()
    torch.cuda.synchronize(device)
    ms = torch.tensor([s.elapsed_time(e)], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.barrier()
    return float(ms.item())


# per entry point: (family, bound, work(args) in bytes or flops) — SURVEY.md §8(d) accounting
def _work_table(nparams=0):
    def dw_fwd(a):   # (x, wt, b, xhat, mean, rstd, B, H, W, C, eps, stream): 2*E*s
        return 2.0 * a[6] * a[7] * a[8] * a[9] * 2
    def dw_bwd(a):   # (dxhat, xhat, rstd, x, dy, wt, dx, dwt, dbias, dycs, B, H, W, C, ...): 3*E*s
        return 3.0 * a[10] * a[11] * a[12] * a[13] * 2     # dy_LN, x in + dx out
    def mlp(a, mi, ci, mult=1.0):
        return mult * 8.0 * a[mi] * a[ci] * a[ci]           # 2 * M * 4C * C
    return {
        "cnx_dwconv_ln_fwd": ("dwconv_ln_fwd", "hbm", dw_fwd),
        "cnx_dwconv_ln_bwd": ("dwconv_ln_bwd", "hbm", dw_bwd),
        "cnx_dwconv_ln_bwd_acc": ("dwconv_ln_bwd", "hbm", dw_bwd),
        "cnx_fc1_gelu_fwd": ("mlp_gemm", "tensor", lambda a: mlp(a, 5, 6)),
        "cnx_fc2_residual_fwd": ("mlp_gemm", "tensor", lambda a: mlp(a, 5, 6)),
        "cnx_mlp_bwd_dz": ("mlp_gemm", "tensor", lambda a: mlp(a, 5, 6)),
        "cnx_mlp_bwd_dxhat": ("mlp_gemm", "tensor", lambda a: mlp(a, 3, 4)),
        "cnx_mlp_bwd_wgrad": ("mlp_gemm", "tensor", lambda a: mlp(a, 6, 7, 2.0)),
        # optimizer step: 4 B/param (read g) + 36 B/param (read p,g,m,v,shadow; write p,m,v,shadow)
        "cnx_optim_grad_norm": ("optim_step", "hbm", lambda a: 4.0 * nparams),
        "cnx_optim_adamw_ema": ("optim_step", "hbm", lambda a: 36.0 * nparams),
    }


def profile_kernels(trainer, device):
    """One extra optimizer step with a CUDA-event pair around every library call (after the
    timed region, so the events do not perturb `value`).  -> {family: stats}, instrumented ms."""
    from image_classification_b200 import _lib
    table = _work_table(sum(p.numel() for p in trainer.params))
    recs, open_ev = [], {}

    def hook(name, phase, args):
        if phase == "pre":
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            open_ev[name] = ev
        else:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            recs.append((name, open_ev.pop(name), ev, args))

    # one stream for the instrumented step: with the weight-gradient GEMMs on the side stream two
    # kernels share the SMs and an event pair would time both (block.py, CNX_BWD_SIDE_STREAM)
    side_env = os.environ.get("CNX_BWD_SIDE_STREAM")
    os.environ["CNX_BWD_SIDE_STREAM"] = "0"
    trainer.step_resident(eager=True)             # re-warm after the e2e loop (eager path)
    torch.cuda.synchronize(device)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _lib.call_hook = hook
    try:
        s.record()
        trainer.step_resident(eager=True)
        e.record()
        torch.cuda.synchronize(device)
    finally:
        _lib.call_hook = None
        if side_env is None:
            del os.environ["CNX_BWD_SIDE_STREAM"]
        else:
            os.environ["CNX_BWD_SIDE_STREAM"] = side_env
    step_ms = s.elapsed_time(e)
    fam = {}
    for name, e0, e1, args in recs:
        family, bound, work = table.get(name, (name.replace("cnx_", ""), "hbm", None))
        d = fam.setdefault(family, dict(bound=bound, calls=0, launches=0, ms=0.0, work=0.0))
        d["calls"] += 1
        d["launches"] += _lib.LAUNCHES_PER_CALL.get(name, 1)
        d["ms"] += e0.elapsed_time(e1)
        if work is not None:
            d["work"] += work(args)
    return fam, step_ms


def kernel_table(fam, timed_step_ms, clocks, pk, keep=None):
    """{family: launches, ms, share of the TIMED step, roofline fraction}."""
    kernels = {}
    for name, d in fam.items():
        if keep is not None and name not in keep:
            continue
        k = {"launches_per_step": d["launches"], "ms_per_step": round(d["ms"], 4),
             "share_of_step": round(d["ms"] / timed_step_ms, 4)}
        if d["work"] > 0:
            if d["bound"] == "hbm":
                k.update(bound="hbm", achieved=round(d["work"] / d["ms"] * 1e-6, 1), unit="GB/s",
                         peak=pk["hbm_gbs"])
            else:
                k.update(bound="tensor", achieved=round(d["work"] / d["ms"] * 1e-9, 1),
                         unit="TFLOP/s", peak=pk["bf16_tflops_sustained"])
            k["frac"] = round(k["achieved"] / k["peak"], 4)
        if name in ("dwconv_ln_fwd", "dwconv_ln_bwd") and d["work"] > 0:
            # the 7x7 depthwise conv is bound by the FP32 FMA pipe before HBM:
            # 49 FMAs per element forward (work = 2*E*2 B), 98 backward (work = 3*E*2 B)
            elems = d["work"] / (4.0 if name == "dwconv_ln_fwd" else 6.0)
            fmas = elems * (49.0 if name == "dwconv_ln_fwd" else 98.0)
            sm_hz = ((clocks or {}).get("sm_mhz") or 1965.0) * 1e6
            peak = 128.0 * SM_COUNT * sm_hz
            k["fma_pipe"] = {"achieved_tfma": round(fmas / d["ms"] * 1e-9, 2),
                             "peak_tfma": round(peak * 1e-12, 2),
                             "frac": round(fmas / (d["ms"] * 1e-3) / peak, 4)}
            if name == "dwconv_ln_bwd":
                # what the implementation actually streams: dxhat, xhat (LayerNorm backward
                # recomputes from the saved normalised tensor), x in + dx out = 4*E*s
                k["declared_4Es_GBps"] = round(d["work"] * 4.0 / 3.0 / d["ms"] * 1e-6, 1)
        kernels[name] = k
    return kernels


def run_convnext_config(tag, device, world, rank, local, micro_batch, res, steps, warmup, pk,
                        headline=False, want_e2e=False, sampler=None):
    """Build a trainer for (micro_batch, res), warm up, time `steps` optimizer steps; -> dict."""
    from image_classification_b200 import _lib
    tr = B200Trainer(device, world, rank, micro_batch, res)
    imgs_per_step = ACCUM * micro_batch * world
    for _ in range(max(3, warmup)):
        tr.step_resident()
    torch.cuda.synchronize(device)
    if tr.reducer is not None:
        # the all-reduce must be live: buckets exist, the unused aux heads were found, steps joined
        assert tr.reducer.buckets and tr.reducer.discovered and tr.reducer.reduced_steps >= 3, \
            "gradient all-reduce did not run during warm-up"
        assert len(tr.reducer.dropped) == (0 if os.environ.get("CNX_NO_AUX") else 6), \
            len(tr.reducer.dropped)
    if sampler is not None:
        sampler.mark()
    l0 = _lib.launch_count
    ms = timed_region(tr.step_resident, steps, world, device)
    launches = _lib.launch_count - l0
    clocks = sampler.stop() if (sampler is not None and rank == 0) else None
    out = {"images_per_s": round(imgs_per_step * steps / (ms * 1e-3), 2),
           "ms_per_step": round(ms / steps, 3), "steps": steps, "global_batch": imgs_per_step,
           "micro_batch_per_gpu": micro_batch, "res": res, "gpu_launches": launches,
           "launch_mode": ("2 CUDA-graph replays (fwd + CE + bwd per micro-batch) + mix, optimizer "
                           "launches" if tr.graphed is not None else "every kernel enqueued from Python")}
    if want_e2e:
        tr.run_e2e(2)
        e2e_losses = []
        ms_e2e = timed_region(lambda: e2e_losses.extend(tr.run_e2e(steps)), 1, world, device)
        assert len(e2e_losses) == steps and all(l == l for l in e2e_losses), e2e_losses
        out["e2e"] = {"value": round(imgs_per_step * steps / (ms_e2e * 1e-3), 2), "unit": UNIT,
                      "h2d_bytes_per_step": tr.h2d_bytes_per_step, "d2h_bytes_per_step": 4,
                      "ms_per_step": round(ms_e2e / steps, 3)}
    if world > 1:
        import torch.distributed as dist
        mine = tr.replica_checksums()
        every = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)
        out["replicas_identical"] = all(torch.equal(every[0], c) for c in every)
        out["allreduce_steps_joined"] = tr.reducer.reduced_steps
        out["allreduce_buckets"] = len(tr.reducer.buckets)
    fam, prof_ms = profile_kernels(tr, device)
    keep = None if headline else ("dwconv_ln_fwd", "dwconv_ln_bwd", "mlp_gemm")
    out["kernels"] = kernel_table(fam, ms / steps, clocks, pk, keep)
    out["profiled_step_ms"] = round(prof_ms, 3)
    out["clocks"] = clocks
    out["timeout_flag"] = _lib.load().cnx_debug_timeout_flag()
    tr.close()
    del tr
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    return out


def run_effnet_config(device, world, rank, steps, warmup, batch=48, res=260):
    """BASELINE config 5: EfficientNet-B2 (torchvision's, timm's architecture; timm is absent),
    260x260, batch 48 per GPU (/root/reference/previous/V3-efb2/config.py:10-15), the V3 loop's
    step (train_advanced_v3.0.py:96-113: bf16 autocast forward, hard-label CE, scaled backward,
    unscale, clip 1.0, AdamW, scaler update) through swap_efficientnet + GraphedForwardBackward +
    FusedOptimStep; data-parallel ranks all-reduce inside the captured graph."""
    import torchvision
    from image_classification_b200 import _lib
    from image_classification_b200.efficientnet import swap_efficientnet
    from image_classification_b200.graphed import GraphedForwardBackward
    from image_classification_b200.loss import SoftTargetCrossEntropyB200
    from image_classification_b200.optim_step import FusedOptimStep
    torch.manual_seed(0)
    model = torchvision.models.efficientnet_b2(weights=None, num_classes=NUM_CLASSES).to(device).train()
    for m in model.modules():                  # timm's create_model default: no drop-path
        if isinstance(m, torchvision.ops.StochasticDepth):
            m.p = 0.0
    swapped = swap_efficientnet(model)
    model = model.to(memory_format=torch.channels_last)
    g = torch.Generator().manual_seed(7 + rank)
    x = torch.randn(batch, 3, res, res, generator=g).to(device).contiguous(
        memory_format=torch.channels_last)
    labels = torch.randint(0, NUM_CLASSES, (batch,), generator=g).to(device)
    opt = torch.optim.AdamW(model.parameters(), lr=2e-4, weight_decay=1e-2)     # V3 config.py:17-18
    scaler = torch.amp.GradScaler("cuda")
    reducer = None
    if world > 1:
        from image_classification_b200.ddp import BucketedGradReducer
        reducer = BucketedGradReducer(list(model.parameters()), bucket_bytes=8 << 20)
    gfb = GraphedForwardBackward(model, SoftTargetCrossEntropyB200(0.0), scaler, x, labels,
                                 grad_reducer=reducer)
    fo = FusedOptimStep(model, opt, scaler, None, 1.0)

    def step():
        gfb(x, labels)
        fo.step()

    for _ in range(max(3, warmup)):
        step()
    l0 = _lib.launch_count
    ms = timed_region(step, steps, world, device)
    out = {"images_per_s": round(world * batch * steps / (ms * 1e-3), 2),
           "ms_per_step": round(ms / steps, 3), "steps": steps, "global_batch": world * batch,
           "batch_per_gpu": batch, "res": res, "modules_swapped": swapped,
           "model": "efficientnet_b2 (torchvision build of timm's architecture), 44 classes",
           "step": "one CUDA-graph replay (fwd + CE + scaled bwd" +
                   (" + bucketed NCCL all-reduce" if world > 1 else "") + ") + 2 optimizer launches",
           "gpu_launches_outside_graph": _lib.launch_count - l0,
           "loss": round(float(gfb.loss), 4)}
    if reducer is not None:
        for h in reducer._hooks:
            h.remove()
    return out


def main_b200(args):
    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    os.environ["CNX_STRICT"] = "1"       # a trunk module off the library path raises, never falls back
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)
    from image_classification_b200 import _lib
    _lib.load()
    _lib.check(_lib.load().cnx_check_device(), "cnx_check_device")
    pk = peaks()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    head = run_convnext_config("cfg2", device, world, rank, local, args.micro_batch, args.res,
                               args.steps, args.warmup, pk, headline=True,
                               want_e2e=not args.no_e2e, sampler=sampler)

    configs = {}
    if not args.no_configs and args.res == RES and args.micro_batch == MICRO_BATCH:
        ks = max(3, args.config_steps)
        micro3 = 1024 // (world * ACCUM) if world > 1 else 64
        c3 = run_convnext_config("cfg3", device, world, rank, local, micro3, 224, ks, 3, pk)
        c3["workload"] = (f"BASELINE config 3: global batch 1024 over {world} GPUs, micro-batch {micro3} "
                          f"x accumulation {ACCUM} (strong scaling)" if world > 1 else
                          "BASELINE config 3, the share of ONE of 8 ranks: micro-batch 64 x "
                          "accumulation 2 on 1 GPU (global batch 1024 needs N >= 2)")
        c3["scaling"] = "strong" if world > 1 else "n/a (one rank's share of the 8-GPU split)"
        configs["cfg3"] = c3
        c4 = run_convnext_config("cfg4", device, world, rank, local, 64, 384, ks, 3, pk)
        c4["workload"] = "BASELINE config 4: ConvNeXt-Base 384x384, micro-batch 64 x accumulation 2 per GPU"
        c4["scaling"] = "weak"
        configs["cfg4"] = c4
        c5 = run_effnet_config(device, world, rank, ks, 3)
        c5["workload"] = "BASELINE config 5: EfficientNet-B2 260x260, batch 48 per GPU"
        c5["scaling"] = "weak"
        configs["cfg5"] = c5

    if rank != 0:
        if world > 1:
            torch.distributed.destroy_process_group()
        return 0

    kernels = head["kernels"]
    dom = max((n for n in kernels if "frac" in kernels[n]), key=lambda n: kernels[n]["ms_per_step"])
    traffic, traffic_note = ncu_traffic() if dom == "mlp_gemm" else (None, None)
    roof = {"kernel": dom, "bound": kernels[dom]["bound"], "achieved": kernels[dom]["achieved"],
            "peak": kernels[dom]["peak"], "unit": kernels[dom]["unit"], "frac": kernels[dom]["frac"],
            "traffic": traffic, "traffic_note": traffic_note,
            "peak_source": pk["source"] +
            (" sustained bf16" if kernels[dom]["bound"] == "tensor" else " copy bandwidth"),
            "how": "sum of algorithmic work over this family's launches in one optimizer step / "
                   "sum of their CUDA-event durations (instrumented step after the timed region)"}

    cpu = None
    if not args.no_cpu_baseline and world == 1:
        ips, cms, cores, sample = run_cpu_reference(5, 1, 8, args.res, "convnext_base")
        cpu = {"value": round(ips, 3), "unit": UNIT, "cores": cores, "kind": "port",
               "sample": sample, "ms_per_step": round(cms, 1), "config1": run_cpu_config1()}

    line = {
        "metric": METRIC, "value": head["images_per_s"], "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": head["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": {"workload": f"ConvNeXt-Base 44-class {args.res}x{args.res} bf16 AMP training "
                               f"step, micro-batch {args.micro_batch} x accumulation {ACCUM} per "
                               "GPU " + ("(BASELINE config 2's per-GPU shape on every rank: weak "
                                         "scaling; config 3's strong-scaling split is under `configs`)"
                                         if args.res == 224 else "(BASELINE config 4 shape)"
                                         if args.res == 384 else "(non-BASELINE shape)"),
                   "global_batch": head["global_batch"], "micro_batch_per_gpu": args.micro_batch,
                   "accumulation": ACCUM, "parallelism": f"dp{world}",
                   "step": "mixup/cutmix + fwd + soft-CE + bwd (x2) + unscale + clip 1.0 + "
                           "AdamW + scheduler + EMA",
                   "l2": "per-step working set (activations + 350 MB fp32 weights, >10 GB) "
                         "exceeds the 126 MB L2; no explicit flush"},
        "e2e": head.get("e2e"), "gpu_launches": head["gpu_launches"], "clocks": head["clocks"],
        "roofline": roof, "kernels": kernels, "cpu_baseline": cpu,
        "profiled_step_ms": head["profiled_step_ms"], "timeout_flag": head["timeout_flag"],
        "configs": configs,
    }
    if world > 1:
        line["replicas_identical"] = head["replicas_identical"]
        line["allreduce"] = {"steps_joined": head["allreduce_steps_joined"],
                             "buckets": head["allreduce_buckets"], "payload_mb": 350.4}
    print(json.dumps(line), flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()
    return 0


if __name__ == "__main__":
    _args = parse_args()
    sys.exit(main_reference(_args) if _args.impl == "reference" else main_b200(_args))
