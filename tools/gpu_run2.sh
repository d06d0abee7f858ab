#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_model.py -q -m gpu -p no:cacheprovider > gpurun_out/test_model.log 2>&1
echo "== model exit=$? : $(tail -1 gpurun_out/test_model.log)"
grep -E "^(E  |FAILED)" gpurun_out/test_model.log | head -20
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err
echo "== bench exit=$?"; tail -3 gpurun_out/bench.err; cat gpurun_out/bench.log
