#!/bin/bash
# ncu --set full on one block fwd+bwd at the stage-1 and stage-3 shapes (tools/ncu_target.py).
# Usage: gpu_ncu2.sh <kernel regex> <tag>
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
REGEX="${1:-dwconv7|ln_rows}"
TAG="${2:-prof}"
python tools/ncu_target.py > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:${REGEX}" -o gpurun_out/${TAG} -f python tools/ncu_target.py > gpurun_out/ncu_${TAG}.log 2>&1
echo "ncu exit=$?"; tail -3 gpurun_out/ncu_${TAG}.log
