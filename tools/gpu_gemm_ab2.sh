#!/bin/bash
# same-box A/B of the GEMM kernels: gpurun_in/libcnx_old.so vs the current build
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_block.py tests/test_gpu_fullsize.py -q -m gpu -x -p no:cacheprovider > gpurun_out/test_gemm.log 2>&1
echo "== tests exit=$? : $(tail -1 gpurun_out/test_gemm.log)"
grep -E "^(E  |FAILED)" gpurun_out/test_gemm.log | cut -c1-300 | head -10
for rep in 1 2; do
CNX_LIB_PATH=$PWD/gpurun_in/libcnx_old.so python tools/microbench.py --out gpurun_out/mb_old.json > gpurun_out/mb_old.log 2>&1
python tools/microbench.py --out gpurun_out/mb_new.json > gpurun_out/mb_new.log 2>&1
python - <<'PY'
import json
a = json.load(open("gpurun_out/mb_old.json")); b = json.load(open("gpurun_out/mb_new.json"))
for x, y in zip(a, b):
    if x["kernel"] in ("fc1_gelu_fwd", "mlp_bwd_dz"):
        print(f"stage {x['stage']} {x['kernel']:16s} old {x['ms']*1e3:7.1f} us   new {y['ms']*1e3:7.1f} us")
PY
done
