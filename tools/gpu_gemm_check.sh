cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_block.py -q -m gpu -p no:cacheprovider -x -k "gemm" > gpurun_out/test_gemm.log 2>&1; echo "gemm tests exit=$? $(tail -1 gpurun_out/test_gemm.log)"
timeout 600 python tools/microbench.py > gpurun_out/microbench.log 2>&1; echo "== microbench exit=$?"
grep -E "fc1|dz|fc2|dxhat" gpurun_out/microbench.log | cut -c1-140
