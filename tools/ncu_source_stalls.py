"""Summarise a `ncu --page source --csv` export: python tools/ncu_source_stalls.py src.csv [top]
Per-SASS-line samples with the dominating stall reason, plus totals per stall reason and per opcode."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]; data = rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = collections.Counter(); byop = collections.Counter(); execs = collections.Counter()
lines = []
for r in data:
    if len(r) < len(hdr): continue
    try: n = int(r[ix["# Samples"]] or 0)
    except ValueError: continue
    src = r[ix["Source"]].strip()
    op = src.split()[0] if not src.startswith("@") else src.split()[1]
    ex = int(r[ix["Instructions Executed"]] or 0)
    execs[op.split(".")[0]] += ex
    byop[op.split(".")[0]] += n
    st = {s: int(r[ix[s]] or 0) for s in stalls}
    for s, v in st.items(): tot[s] += v
    lines.append((n, r[ix["Address"]], src, ex, max(st, key=st.get) if n else ""))
N = sum(l[0] for l in lines)
print("total samples", N, " total warp-instr executed", sum(execs.values()))
print("stall totals:", ", ".join(f"{k[6:]}={v} ({100*v/max(N,1):.0f}%)" for k, v in tot.most_common(10)))
print("samples by opcode:", ", ".join(f"{k}={v}" for k, v in byop.most_common(16)))
print("executed by opcode:", ", ".join(f"{k}={v}" for k, v in execs.most_common(24)))
for n, a, s, ex, why in sorted(lines, reverse=True)[:top]:
    print(f"{n:7d} {100*n/max(N,1):5.1f}%  ex={ex:9d} {why[6:]:14s} {s[:90]}")
