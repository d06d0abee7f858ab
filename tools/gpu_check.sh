#!/bin/bash
# Whole -m gpu suite, smoke, default bench on ONE GPU (no reference arm).
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/test_all.log 2>&1
echo "== gpu tests exit=$? : $(tail -1 gpurun_out/test_all.log)"
grep -E "^(E  |FAILED)" gpurun_out/test_all.log | cut -c1-300 | head -20
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "== smoke exit=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err
echo "== bench exit=$?"; tail -5 gpurun_out/bench.err; cut -c1-1500 gpurun_out/bench.log
