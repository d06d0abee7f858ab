#!/bin/bash
# Usage: gpu_quick.sh "<pytest -k expr>" [bench args...]: selected GPU tests, then bench.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
K="$1"; shift
timeout 900 python -m pytest tests -q -m gpu -x -p no:cacheprovider -k "$K" > gpurun_out/test_sel.log 2>&1
echo "== tests exit=$? : $(tail -1 gpurun_out/test_sel.log)"
grep -E "^(E  |FAILED)" gpurun_out/test_sel.log | cut -c1-400 | head -20
timeout 900 python bench.py "$@" > gpurun_out/bench.log 2> gpurun_out/bench.err
echo "== bench exit=$?"; tail -5 gpurun_out/bench.err; cut -c1-400 gpurun_out/bench.log
