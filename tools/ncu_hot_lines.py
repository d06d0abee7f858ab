"""Top stall lines (deduplicated SASS) of one launch in an .ncu-rep:
python tools/ncu_hot_lines.py rep.ncu-rep <kernel regex> <launch index> [min samples]"""
import csv, io, subprocess, sys
rep, rx, idx = sys.argv[1], sys.argv[2], int(sys.argv[3])
mins = int(sys.argv[4]) if len(sys.argv) > 4 else 20
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{rx}",
                      "--launch-skip", str(idx), "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
print(rows[0][1][:100])
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
seen, lines = set(), []
for r in rows[2:]:
    if len(r) < len(hdr) or r[ix["Address"]] in seen: continue
    seen.add(r[ix["Address"]])
    try: n = int(r[ix["# Samples"]] or 0)
    except ValueError: continue
    st = {s[6:]: int(r[ix[s]] or 0) for s in stalls}
    lines.append((r[ix["Address"]], n, int(r[ix["Instructions Executed"]] or 0), r[ix["Source"]].strip(), st))
N = sum(l[1] for l in lines)
print("samples", N)
for a, n, ex, src, st in lines:
    if n >= mins:
        top = ",".join(f"{k}:{v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:2] if v)
        print(f"{a[-5:]} {n:5d} {100*n/N:5.1f}% ex={ex:8d} {src[:72]:72s} {top}")
