#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 9000 -c 4000 --csv --log-file gpurun_out/launches_r01.csv $CMD > gpurun_out/ncu_bench.log 2>&1
echo "launch list exit=$? lines=$(wc -l < gpurun_out/launches_r01.csv)"
python tools/ncu_target.py > gpurun_out/ncu_plain_target.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"dwconv_ln|gemm_bf16" -s 8 -c 8 -o gpurun_out/prof_r01_s1 python tools/ncu_target.py > gpurun_out/ncu_target1.log 2>&1
echo "full s1 exit=$?"
ncu --set full --clock-control none --import-source on -k regex:"dwconv_ln|gemm_bf16" -s 24 -c 8 -o gpurun_out/prof_r01_s3 python tools/ncu_target.py > gpurun_out/ncu_target3.log 2>&1
echo "full s3 exit=$?"; ls -la gpurun_out/*.ncu-rep
