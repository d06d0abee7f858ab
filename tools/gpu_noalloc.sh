#!/bin/bash
# The allocation-free entry point's test alone in a fresh process, then tools/gpu_check.sh (whole
# suite with that file collected last, smoke, default bench).
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider > gpurun_out/test_noalloc_alone.log 2>&1
echo "== no-alloc test alone exit=$? : $(tail -1 gpurun_out/test_noalloc_alone.log)"
grep -E "^(E  |FAILED)" gpurun_out/test_noalloc_alone.log | cut -c1-300 | head -20
bash tools/gpu_check.sh
