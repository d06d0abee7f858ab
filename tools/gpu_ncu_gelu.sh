#!/bin/bash
# ncu --set full with source-level sampling of the two GELU-carrying GEMMs at the stage-3 shape.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
S=${1:-2}; EW=${2:-0}
timeout 300 python tools/ncu_gelu_target.py $S $EW > gpurun_out/ncu_gelu_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:gemm2" --launch-skip 4 --launch-count 2 \
  -o gpurun_out/prof_gelu_s${S}_ew${EW} -f python tools/ncu_gelu_target.py $S $EW > gpurun_out/ncu_gelu.log 2>&1
echo "exit=$?"; tail -3 gpurun_out/ncu_gelu.log; ls -la gpurun_out/*.ncu-rep
