#!/bin/bash
# Full GPU pass: whole -m gpu suite, smoke, microbench, bench.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -p no:cacheprovider -x > gpurun_out/test_all.log 2>&1
echo "== gpu tests exit=$? : $(tail -1 gpurun_out/test_all.log)"
grep -E "^(E  |FAILED)" gpurun_out/test_all.log | cut -c1-300 | head -20
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "== smoke exit=$?: $(tail -1 gpurun_out/smoke.log | cut -c1-300)"
timeout 600 python tools/microbench.py > gpurun_out/microbench.log 2>&1; echo "== microbench exit=$?"
grep -E "wgrad|dxhat|fc2" gpurun_out/microbench.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err
echo "== bench exit=$?"; tail -3 gpurun_out/bench.err; cat gpurun_out/bench.log
