#!/bin/bash
# ncu launch list (gpu__time_duration only) of one bench optimizer step. Usage: gpu_launchlist.sh <tag> [skip]
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
TAG="${1:-r01}"
SKIP="${2:-11000}"
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s $SKIP -c 5000 --csv --log-file gpurun_out/launches_${TAG}.csv $CMD > gpurun_out/ncu_bench.log 2>&1
echo "launch list exit=$? lines=$(wc -l < gpurun_out/launches_${TAG}.csv)"
