#!/bin/bash
# Round-2 evidence: (1) launch list (device time of every launch) of one eager optimizer step of the
# bench; (2) ncu --set full of one block forward + backward at the stage-1 and stage-3 shapes.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
export CNX_NO_GRAPH=1     # the launch list wants one ncu record per kernel launch enqueued from Python
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-configs"
$CMD > gpurun_out/ncu_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file gpurun_out/launches_r02.csv $CMD > gpurun_out/ncu_bench.log 2>&1
echo "launch list exit=$? lines=$(wc -l < gpurun_out/launches_r02.csv)"
python tools/ncu_target.py > gpurun_out/ncu_plain_target.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:dwconv7|ln_rows|gemm2" -o gpurun_out/prof_r02 -f python tools/ncu_target.py > gpurun_out/ncu_target.log 2>&1
echo "full exit=$?"; ls -la gpurun_out/*.ncu-rep; tail -2 gpurun_out/ncu_target.log
