#!/bin/bash
# ncu --set full with source-level sampling of the conv kernels at one stage shape; summarised here
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
S=${1:-2}
timeout 200 python tools/ncu_conv_target.py $S > gpurun_out/ncu_conv_plain.log 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k "regex:dwconv7" --launch-skip 4 --launch-count 2 \
  -o gpurun_out/prof_conv_s${S} -f python tools/ncu_conv_target.py $S > gpurun_out/ncu_conv.log 2>&1
echo "exit=$?"; tail -2 gpurun_out/ncu_conv.log; ls -la gpurun_out/*.ncu-rep
