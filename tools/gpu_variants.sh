#!/bin/bash
# Development: time the MLP GEMMs with every gpurun_in/libcnx_*.so variant (and the current build).
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
ARGS="$@"
for rep in 1 2; do
python tools/gemm_quick.py --tag current $ARGS 2>&1 | tail -1
for lib in gpurun_in/libcnx_*.so; do
  CNX_LIB_PATH=$PWD/$lib python tools/gemm_quick.py --tag $(basename $lib .so) $ARGS 2>&1 | tail -1
done
done | tee gpurun_out/variants.log
