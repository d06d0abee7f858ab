"""Where the gradient all-reduce sits in the step: kernel timeline of ONE
optimizer step of bench.py's trainer on N GPUs, from torch.profiler (CUPTI), summarised as text.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29613 tools/nccl_overlap.py [--micro-batch 128] [--out profiles/...txt]

For every NCCL kernel of the profiled step: start (ms after the step's first kernel), duration, and
how much of it ran while one of the library's own kernels (cnx::*) was executing on another stream.
Rank 0 writes the summary.  Development / evidence tool; nothing here is timed for the bench.
"""
from __future__ import annotations

import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("CNX_STRICT", "1")

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--micro-batch", type=int, default=128)
    ap.add_argument("--res", type=int, default=224)
    ap.add_argument("--out", default="gpurun_out/nccl_overlap.txt")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    import bench
    tr = bench.B200Trainer(dev, world, rank, args.micro_batch, args.res)
    for _ in range(4):
        tr.step_resident()
    torch.cuda.synchronize()
    dist.barrier()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        tr.step_resident()
        torch.cuda.synchronize()
    if rank != 0:            # rank 0 alone reads the trace; no collective after the profiled step
        os._exit(0)          # (a barrier here hung until the caller's timeout on the 2-GPU box)
    ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA
          and e.time_range is not None]
    kern = []
    for e in ev:
        name = e.name
        if name.startswith("Memcpy") or name.startswith("Memset"):
            continue
        kern.append((e.time_range.start, e.time_range.end, name))
    kern.sort()
    if rank == 0 and kern:
        t0 = kern[0][0]
        t1 = max(k[1] for k in kern)
        own = [(s, e) for s, e, n in kern if "cnx" in n]
        nccl = [(s, e, n) for s, e, n in kern if "nccl" in n.lower()]
        # union of the library's kernel intervals
        merged = []
        for s, e in sorted(own):
            if merged and s <= merged[-1][1]:
                merged[-1][1] = max(merged[-1][1], e)
            else:
                merged.append([s, e])

        def overlap(s, e):
            tot = 0.0
            for a, b in merged:
                if b <= s:
                    continue
                if a >= e:
                    break
                tot += min(b, e) - max(a, s)
            return tot
        bwd_last = max(e for s, e in own)
        lines = [f"world {world}, micro-batch {args.micro_batch} x 2, {args.res}^2; one optimizer step "
                 f"(2 CUDA-graph replays + optimizer), rank 0; times in ms after the first kernel",
                 f"step span {(t1 - t0) / 1e3:.2f} ms, {len(kern)} kernels, {len(own)} cnx:: kernels "
                 f"({sum(e - s for s, e in own) / 1e3:.2f} ms summed), {len(nccl)} NCCL kernels "
                 f"({sum(e - s for s, e, _ in nccl) / 1e3:.2f} ms summed)",
                 f"last cnx:: kernel ends at {(bwd_last - t0) / 1e3:.2f} ms",
                 "NCCL kernel                              start    dur   under cnx:: kernels"]
        tot_d = tot_o = 0.0
        for s, e, n in nccl:
            o = overlap(s, e)
            tot_d += e - s
            tot_o += o
            lines.append(f"{n[:38]:38s} {(s - t0) / 1e3:7.2f} {(e - s) / 1e3:6.3f}   {100 * o / max(e - s, 1e-9):5.1f} %")
        lines.append(f"all-reduce time overlapped with the library's kernels: {tot_o / 1e3:.3f} of "
                     f"{tot_d / 1e3:.3f} ms ({100 * tot_o / max(tot_d, 1e-9):.1f} %)")
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        with open(args.out, "w") as f:
            f.write("\n".join(lines) + "\n")
        print("\n".join(lines), flush=True)
    os._exit(0)


if __name__ == "__main__":
    main()
