"""Why is the host-fed (e2e) loop slower than the resident one?  Same process, auxiliary heads on /
off, 5 and 20 steps, with the caching allocator's cudaMalloc counters.  `python tools/e2e_diag.py`"""
import json, os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench

dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
tr = bench.B200Trainer(dev, 1, 0, 128, 224)
for _ in range(3):
    tr.step_resident()
tr.run_e2e(2)
torch.cuda.synchronize()


def stat(k):
    return torch.cuda.memory_stats(dev).get(k, 0)


for aux in (True, False, True, False):
    tr.model.compute_aux = aux
    tr.step_resident(); tr.run_e2e(2); torch.cuda.synchronize()
    r = {"aux": aux}
    r["resident_ms"] = round(bench.timed_region(tr.step_resident, 5, 1, dev) / 5, 3)
    for n in (5, 20):
        a0, s0 = stat("num_device_alloc"), stat("num_sync_all_streams")
        t0 = time.perf_counter()
        ms = bench.timed_region(lambda: tr.run_e2e(n), 1, 1, dev)
        r[f"e2e{n}_ms"] = round(ms / n, 3)
        r[f"e2e{n}_wall_ms"] = round((time.perf_counter() - t0) * 1e3 / n, 3)
        r[f"e2e{n}_cudaMalloc"] = stat("num_device_alloc") - a0
        r[f"e2e{n}_sync_all"] = stat("num_sync_all_streams") - s0
    print(json.dumps(r), flush=True)
