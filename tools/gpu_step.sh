#!/bin/bash
# One GPU visit: parity tests for the kernels, then per-kernel timings.  Usage: gpu_step.sh [pytest -k expr]
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
KEXPR="${1:-}"
if [ -n "$KEXPR" ]; then
  timeout 900 python -m pytest tests/test_gpu_block.py -q -m gpu -p no:cacheprovider -x -k "$KEXPR" > gpurun_out/test_block.log 2>&1
else
  timeout 900 python -m pytest tests/test_gpu_block.py -q -m gpu -p no:cacheprovider > gpurun_out/test_block.log 2>&1
fi
echo "== block tests exit=$? : $(tail -1 gpurun_out/test_block.log)"
grep -E "^(E  |FAILED)" gpurun_out/test_block.log | head -30
timeout 600 python tools/microbench.py > gpurun_out/microbench.log 2>&1
echo "== microbench exit=$?"
cat gpurun_out/microbench.log | tail -45
