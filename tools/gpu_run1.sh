#!/bin/bash
# First GPU pass.  Every kernel family first runs once under compute-sanitizer memcheck (an
# out-of-bounds access is then reported by the tool instead of faulting the GPU); only families
# that come out clean go on to the plain parity tests (separate processes per group).
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
CS="compute-sanitizer --tool memcheck --error-exitcode 9 --print-limit 20"
san() { # family
  timeout 600 $CS python tools/sanity_kernels.py $1 > gpurun_out/san_$1.log 2>&1
  local rc=$?
  echo "== sanitizer $1 exit=$rc : $(grep -E 'ERROR SUMMARY|rel err|timeout_flag=1' gpurun_out/san_$1.log | tr '\n' ' ' | cut -c1-400)"
  return $rc
}
run() { # name, pytest args...
  local name=$1; shift
  timeout 600 python -m pytest "$@" -q -m gpu -p no:cacheprovider > gpurun_out/test_$name.log 2>&1
  echo "== $name exit=$? : $(tail -1 gpurun_out/test_$name.log)"
}
san glue && run glue tests/test_gpu_glue.py
if san gemm; then
  run gemm_plain tests/test_gpu_block.py -k gemm_plain
  run gemm_mn tests/test_gpu_block.py -k gemm_mn
  run gemm_epi tests/test_gpu_block.py -k gemm_epilogues
  GEMM_OK=1
fi
if san dw; then
  run dwfwd tests/test_gpu_block.py -k dwconv_ln_fwd
  run dwbwd tests/test_gpu_block.py -k dwconv_ln_bwd
  DW_OK=1
fi
if [ -n "$GEMM_OK" ] && [ -n "$DW_OK" ] && san block; then
  run block tests/test_gpu_block.py -k "block or fold"
  timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "== smoke exit=$?: $(tail -1 gpurun_out/smoke.log)"
  timeout 900 python tools/microbench.py > gpurun_out/microbench.log 2>&1; echo "== microbench exit=$?"; tail -45 gpurun_out/microbench.log
fi
