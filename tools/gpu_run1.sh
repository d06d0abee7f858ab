#!/bin/bash
# GPU pass: smallest launches first (tools/sanity_kernels.py, one process per kernel family),
# then the parity tests group by group; a family that fails its sanity run is not tested further.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
san() { # family
  timeout 300 python tools/sanity_kernels.py $1 > gpurun_out/san_$1.log 2>&1
  local rc=$?
  echo "== sanity $1 exit=$rc : $(grep -E 'rel err|timeout_flag=1|Error|error' gpurun_out/san_$1.log | tr '\n' ' ' | cut -c1-600)"
  return $rc
}
run() { # name, pytest args...
  local name=$1; shift
  timeout 600 python -m pytest "$@" -q -m gpu -p no:cacheprovider > gpurun_out/test_$name.log 2>&1
  echo "== $name exit=$? : $(tail -1 gpurun_out/test_$name.log)"
}
FAM="${1:-all}"
if [ "$FAM" = all ] || [ "$FAM" = glue ]; then san glue && run glue tests/test_gpu_glue.py; fi
if [ "$FAM" = all ] || [ "$FAM" = gemm ]; then
  if san gemm; then
    run gemm_plain tests/test_gpu_block.py -k gemm_plain
    run gemm_mn tests/test_gpu_block.py -k gemm_mn
    run gemm_epi tests/test_gpu_block.py -k gemm_epilogues
  fi
fi
if [ "$FAM" = all ] || [ "$FAM" = dw ]; then
  if san dw; then
    run dwfwd tests/test_gpu_block.py -k dwconv_ln_fwd
    run dwbwd tests/test_gpu_block.py -k dwconv_ln_bwd
  fi
fi
if [ "$FAM" = all ] || [ "$FAM" = block ]; then
  if san block; then
    run block tests/test_gpu_block.py -k "block or fold"
    timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "== smoke exit=$?: $(tail -1 gpurun_out/smoke.log)"
    timeout 900 python tools/microbench.py > gpurun_out/microbench.log 2>&1; echo "== microbench exit=$?"; tail -45 gpurun_out/microbench.log
  fi
fi
nvidia-smi -q | grep -i -A2 "xid" | head; dmesg 2>/dev/null | grep -i xid | tail -3
