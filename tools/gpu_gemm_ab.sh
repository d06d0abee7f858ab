#!/bin/bash
# GEMM epilogue-warp A/B: parity of the forced instantiations, then the per-kernel microbench with
# 8, 16 and automatic epilogue warps.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_block.py -q -m gpu -x -p no:cacheprovider -k "gemm" > gpurun_out/test_gemm.log 2>&1
echo "== gemm tests exit=$? : $(tail -1 gpurun_out/test_gemm.log)"
grep -E "^(E  |FAILED)" gpurun_out/test_gemm.log | cut -c1-300 | head -10
for ew in 8 16 0; do
  timeout 600 python tools/microbench.py --epi-warps $ew --out gpurun_out/microbench_ew$ew.json > gpurun_out/microbench_ew$ew.log 2>&1
  echo "== microbench ew=$ew exit=$?"
done
python - <<'PY'
import json
rows = {ew: json.load(open(f"gpurun_out/microbench_ew{ew}.json")) for ew in (8, 16, 0)}
for i, r in enumerate(rows[8]):
    if "cublas" in r["kernel"] or "dwconv" in r["kernel"]:
        continue
    print(f"stage {r['stage']} {r['kernel']:22s} ew8 {r['ms']*1e3:7.1f} us  ew16 {rows[16][i]['ms']*1e3:7.1f} us  auto {rows[0][i]['ms']*1e3:7.1f} us")
PY
