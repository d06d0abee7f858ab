#!/bin/bash
# In-step A/B of an environment switch: bash tools/gpu_step_ab.sh VAR [steps]
# runs the default 1-GPU bench with VAR=1 and without, twice each, and prints ms/step.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
VAR=$1; STEPS=${2:-10}
timeout 900 python -m pytest tests/test_gpu_block.py -q -m gpu -x -p no:cacheprovider > gpurun_out/test_gemm.log 2>&1
echo "== tests exit=$? : $(tail -1 gpurun_out/test_gemm.log)"
grep -E "^(E  |FAILED)" gpurun_out/test_gemm.log | cut -c1-300 | head -10
for rep in 1 2; do
for on in 1 0; do
  if [ $on = 1 ]; then export $VAR=1; else unset $VAR; fi
  timeout 600 python bench.py --steps $STEPS --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/ab_$on.json 2> gpurun_out/ab_$on.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/ab_$on.json").read().strip().splitlines()[-1])
k = d.get("kernels", {})
print("$VAR=" + ("1" if $on else "unset"), "ms/step", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"],
      "mlp_gemm", k.get("mlp_gemm", {}).get("ms_per_step"), "frac", d["roofline"]["frac"],
      "dw_fwd", k.get("dwconv_ln_fwd", {}).get("ms_per_step"), "dw_bwd", k.get("dwconv_ln_bwd", {}).get("ms_per_step"))
PY
done
done | tee gpurun_out/step_ab.log
