#!/bin/bash
# 32- vs 64-column staging chunks (one more operand stage): parity tests, then A/B timings
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_block.py tests/test_gpu_fullsize.py -q -m gpu -x -p no:cacheprovider > gpurun_out/test_gemm.log 2>&1
echo "== tests exit=$? : $(tail -1 gpurun_out/test_gemm.log)"
grep -E "^(E  |FAILED)" gpurun_out/test_gemm.log | cut -c1-300 | head -10
for rep in 1 2; do
CNX_GEMM_WIDE_CHUNKS=1 timeout 300 python tools/gemm_quick.py --tag wide64 2>&1 | tail -1
timeout 300 python tools/gemm_quick.py --tag narrow32 2>&1 | tail -1
done | tee gpurun_out/chunks_ab.log
