"""Dump deduplicated SASS lines of one `ncu --page source --csv` export with offsets from the kernel
start: python tools/ncu_lines.py src.csv [lo_hex hi_hex] [min_samples]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
lo = int(sys.argv[2], 16) if len(sys.argv) > 2 else 0
hi = int(sys.argv[3], 16) if len(sys.argv) > 3 else 1 << 30
mins = int(sys.argv[4]) if len(sys.argv) > 4 else 0
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
seen, out = set(), []
for r in rows[2:]:
    if len(r) < len(hdr) or r[ix["Address"]] in seen or r[ix["Address"]] == "Address": continue
    seen.add(r[ix["Address"]]); out.append(r)
base = int(out[0][ix["Address"]], 16)
tot = 0
for r in out:
    a = int(r[ix["Address"]], 16) - base
    n = int(r[ix["# Samples"]] or 0)
    if lo <= a <= hi:
        tot += n
        if n >= mins:
            st = {s[6:]: int(r[ix[s]] or 0) for s in stalls}
            top = ",".join(f"{k}:{v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:2] if v)
            print(f"{a:5x} {n:5d} ex={int(r[ix['Instructions Executed']] or 0):8d} {r[ix['Source']].strip()[:80]:80s} {top}")
print("samples in range", tot)
