"""Is the step launch-bound?  CPU time to ENQUEUE one optimizer step (no sync inside) next to the
GPU time of the same step, plus the forward / backward / optimizer split of both."""
import os, sys, time, json, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench

dev = torch.device("cuda:0")
tr = bench.B200Trainer(dev, 1, 0, 128, 224)
for _ in range(4):
    tr.step_resident()
torch.cuda.synchronize()
res = []
for _ in range(5):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    tr.step_resident()
    e1.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    res.append(dict(cpu_enqueue_ms=round((t1 - t0) * 1e3, 2), gpu_ms=round(e0.elapsed_time(e1), 2),
                    wall_ms=round((t2 - t0) * 1e3, 2)))
print(json.dumps(res))

# phase split: sync between phases of one micro-step (changes overlap, shows each phase's floor)
model, crit = tr.model, tr.criterion
x = tr.dev_images[0].contiguous(memory_format=torch.channels_last)
tg = torch.nn.functional.one_hot(tr.dev_labels[0], 44).float()
def timed(fn):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); out = fn(); e1.record(); t1 = time.perf_counter(); torch.cuda.synchronize()
    return out, round((t1 - t0) * 1e3, 2), round(e0.elapsed_time(e1), 2)
for _ in range(3):
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out, cf, gf = timed(lambda: crit(model(x), tg))
    _, cb, gb = timed(lambda: out.backward())
    print(json.dumps(dict(fwd_cpu_ms=cf, fwd_gpu_ms=gf, bwd_cpu_ms=cb, bwd_gpu_ms=gb)))
