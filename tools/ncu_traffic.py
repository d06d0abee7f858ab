"""profiles/r02_ncu_gemm_traffic.json from an `ncu --set full` capture of tools/ncu_target.py:
DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per launch of the MLP GEMM kernels of
one stage-3 block (C=512, M=25088: the shape 27 of the 36 blocks run), next to their algorithmic
bytes, and the same for the dwconv / LayerNorm kernels of the stage-1 block (fused forward).
bench.py reads `bytes_per_launch_mean` for roofline.traffic.

    python tools/ncu_traffic.py gpurun_out/prof_r02.ncu-rep <git-hash> [out.json]
"""
import csv
import io
import json
import re
import subprocess
import sys

rep, git = sys.argv[1], sys.argv[2]
out_path = sys.argv[3] if len(sys.argv) > 3 else "profiles/r02_ncu_gemm_traffic.json"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}


def val(r, name):
    v = float(r[idx[name]].replace(",", ""))
    u = units[idx[name]].lower()
    mult = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1e-3, "nsecond": 1e-3,
            "us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3}.get(u, 1)
    return v * mult


launches = []
for r in data:
    name = re.sub(r"\(.*", "", r[idx["Kernel Name"]])
    name = re.sub(r"^.*::", "", name)
    launches.append({"kernel": name, "grid": r[idx["Grid Size"]] if "Grid Size" in idx else "",
                     "us": round(val(r, "gpu__time_duration.sum"), 2),
                     "dram_read": val(r, "dram__bytes_read.sum"), "dram_write": val(r, "dram__bytes_write.sum"),
                     "tensor_pct": float(r[idx["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]] or 0)
                     if "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active" in idx else None})
# tools/ncu_target.py: stage-1 block (C=128, H=56) warm-up pass + profiled pass, then stage 3
# (C=512, H=14) likewise; the capture holds every matching launch: the LAST pass of each config
gemm = [l for l in launches if l["kernel"].startswith("gemm2_kernel")]
n = len(gemm) // 4                       # 6 GEMM launches per pass, 2 passes x 2 configs
s3 = gemm[3 * n:]
M, C = 25088, 512
alg = {"fc1": 2 * M * C + 2 * 4 * C * C + 2 * 2 * M * 4 * C, "fc2": 2 * M * 4 * C + 2 * 4 * C * C + 2 * 2 * M * C,
       "dz": 2 * M * C + 2 * 4 * C * C + 2 * 2 * M * 4 * C, "dxhat": 2 * M * 4 * C + 2 * 4 * C * C + 2 * M * C,
       "wgrad": 2 * M * 4 * C + 2 * M * C + 4 * 4 * C * C}
alg_mean = (alg["fc1"] + alg["fc2"] + alg["dz"] + alg["dxhat"] + 2 * alg["wgrad"]) / 6.0
doc = {
    "git": git,
    "what": "dram__bytes_read.sum + dram__bytes_write.sum per launch, mean of the 6 MLP GEMM launches "
            "of one stage-3 block (C=512, M=25088), ncu --set full --clock-control none on tools/ncu_target.py",
    "bytes_per_launch_mean": round(sum(l["dram_read"] + l["dram_write"] for l in s3) / max(1, len(s3))),
    "algorithmic_bytes_mean": round(alg_mean),
    "stage3_gemm_launches": s3,
    "all_launches": launches,
}
E1 = 128 * 56 * 56 * 128
for l in launches:
    if l["kernel"].startswith("dwconv7_fwd") and "stage1_fwd" not in doc and l["dram_read"] > 0.5 * E1 * 2:
        doc["stage1_fwd"] = l
for l in launches:
    if l["kernel"].startswith("dwconv7_fwd") and l["dram_read"] > 0.5 * E1 * 2:
        doc["stage1_dwconv_ln_fwd"] = {"launch": l, "algorithmic_bytes": 2 * E1 * 2,
                                       "dram_over_algorithmic": round((l["dram_read"] + l["dram_write"]) / (2 * E1 * 2), 3)}
doc.pop("stage1_fwd", None)
json.dump(doc, open(out_path, "w"), indent=1)
print(out_path, doc["bytes_per_launch_mean"], doc["algorithmic_bytes_mean"], doc.get("stage1_dwconv_ln_fwd", {}).get("dram_over_algorithmic"))
