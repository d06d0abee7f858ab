#!/bin/bash
# Round-2 final records on ONE GPU: default bench, then the ncu evidence of the same build
# (launch list of one eager step; --set full of one block at stage 1 and 3), summarised ON the box
# (the .ncu-rep is deleted: gpurun_out/ must stay under 64 MiB to travel back).
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
GIT=${1:-unknown}
timeout 900 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err
echo "== bench exit=$?"; tail -3 gpurun_out/bench.err; cut -c1-400 gpurun_out/bench.log
export CNX_NO_GRAPH=1
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-configs"
timeout 300 $CMD > gpurun_out/ncu_plain_bench.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file gpurun_out/launches_r02.csv $CMD > gpurun_out/ncu_bench.log 2>&1
echo "launch list exit=$? lines=$(wc -l < gpurun_out/launches_r02.csv)"
python tools/launch_summary.py gpurun_out/launches_r02.csv 40 > gpurun_out/launches_summary.txt 2>&1; head -12 gpurun_out/launches_summary.txt
rm -f gpurun_out/launches_r02.csv
timeout 300 python tools/ncu_target.py > gpurun_out/ncu_plain_target.log 2>&1 &&
timeout 600 ncu --set full --clock-control none -k "regex:dwconv7|ln_rows|gemm2" -o gpurun_out/prof_r02 -f python tools/ncu_target.py > gpurun_out/ncu_target.log 2>&1
echo "full exit=$?"; ls -la gpurun_out/*.ncu-rep
python tools/ncu_summary.py gpurun_out/prof_r02.ncu-rep > gpurun_out/ncu_block_summary.txt 2>&1
python tools/ncu_traffic.py gpurun_out/prof_r02.ncu-rep $GIT gpurun_out/ncu_gemm_traffic.json > gpurun_out/ncu_traffic.log 2>&1; tail -3 gpurun_out/ncu_traffic.log
rm -f gpurun_out/*.ncu-rep
du -sh gpurun_out
