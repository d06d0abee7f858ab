"""Summarise an ncu launch list (csv with gpu__time_duration.sum) over ONE optimizer step.
A step = 2 micro-batches, each opened by mix_batch_kernel: the window runs from the 4th-last to the
2nd-last mix_batch launch.  python tools/launch_summary.py launches.csv"""
import csv, re, sys, collections
rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    us = v / 1000.0 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1000.0)
    rows.append((r["Kernel Name"], us))
mix = [i for i, (n, _) in enumerate(rows) if "mix_batch" in n]
if len(mix) >= 4:
    lo, hi = mix[-4], mix[-2]
else:
    lo, hi = 0, len(rows)
win = rows[lo:hi]
def short(n):
    n = re.sub(r"^void ", "", n)
    n = re.sub(r"\(.*", "", n)
    n = n.replace("cnx::<unnamed>::", "").replace("cnx::", "")
    if n.startswith("at::") or "cutlass" in n or "cudnn" in n or "nchw" in n.lower() or "elementwise" in n:
        n = "torch/other: " + n[:70]
    return n[:90]
agg = collections.OrderedDict()
for n, us in win:
    k = short(n)
    a = agg.setdefault(k, [0.0, 0])
    a[0] += us; a[1] += 1
tot = sum(a[0] for a in agg.values())
print(f"window = launches [{lo},{hi}) of {len(rows)} captured: one optimizer step; total {tot:.1f} us over {len(win)} launches")
for k, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:int(sys.argv[2]) if len(sys.argv) > 2 else 40]:
    print(f"{us:10.1f} us {100*us/tot:5.1f}% n={n:4d} avg={us/n:8.1f} us  {k}")
