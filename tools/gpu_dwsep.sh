#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_dwsep.py -q -m gpu -p no:cacheprovider > gpurun_out/test_dwsep.log 2>&1; echo "exit=$?"; grep -E "^(E  |FAILED)" gpurun_out/test_dwsep.log | cut -c1-250 | head -40; tail -3 gpurun_out/test_dwsep.log
timeout 600 python tools/effnet_bench.py > gpurun_out/effnet_bench.json 2> gpurun_out/effnet_bench.err; echo "bench exit=$?"; tail -c 3000 gpurun_out/effnet_bench.json; tail -5 gpurun_out/effnet_bench.err
