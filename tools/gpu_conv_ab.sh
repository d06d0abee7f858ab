#!/bin/bash
# Same-box A/B of the conv kernels: gpurun_in/libcnx_old.so (previous build) vs the current library
# (fused LayerNorm epilogue for C <= 128) vs the current library with CNX_DWCONV_NO_FUSE=1.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_block.py tests/test_gpu_fullsize.py -q -m gpu -x -p no:cacheprovider -k "dwconv or block or full" > gpurun_out/test_conv.log 2>&1
echo "== conv tests exit=$? : $(tail -1 gpurun_out/test_conv.log)"
for rep in 1 2; do
CNX_LIB_PATH=$PWD/gpurun_in/libcnx_old.so python tools/microbench.py --stages 0,1 --out gpurun_out/mb_old.json > gpurun_out/mb_old.log 2>&1
python tools/microbench.py --stages 0,1 --out gpurun_out/mb_new.json > gpurun_out/mb_new.log 2>&1
CNX_DWCONV_NO_FUSE=1 python tools/microbench.py --stages 0,1 --out gpurun_out/mb_nofuse.json > gpurun_out/mb_nofuse.log 2>&1
python - <<'PY'
import json
a = json.load(open("gpurun_out/mb_old.json")); b = json.load(open("gpurun_out/mb_new.json")); c = json.load(open("gpurun_out/mb_nofuse.json"))
for x, y, z in zip(a, b, c):
    if "dwconv" in x["kernel"]:
        print(f"stage {x['stage']} {x['kernel']:16s} old {x['ms']*1e3:7.1f} us   new {y['ms']*1e3:7.1f} us   new, CNX_DWCONV_NO_FUSE {z['ms']*1e3:7.1f} us")
PY
done
