#!/bin/bash
# dwsep GPU tests, the effnet bench, then ncu --set full of the BatchNorm / squeeze-excite kernels.
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_dwsep.py -q -m gpu -p no:cacheprovider > gpurun_out/test_dwsep.log 2>&1; echo "exit=$?"; grep -E "^(E  |FAILED)" gpurun_out/test_dwsep.log | cut -c1-250 | head -20; tail -2 gpurun_out/test_dwsep.log
timeout 600 python tools/effnet_bench.py > gpurun_out/effnet_bench.json 2> gpurun_out/effnet_bench.err; echo "bench exit=$?"
timeout 300 python tools/effnet_bench.py --ncu > gpurun_out/ncu_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:bn_|se_" -o gpurun_out/r01_bnse -f python tools/effnet_bench.py --ncu > gpurun_out/ncu_bnse.log 2>&1
echo "ncu exit=$?"; tail -3 gpurun_out/ncu_bnse.log
