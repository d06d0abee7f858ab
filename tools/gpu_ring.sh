#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
export CNX_GEMM_NO_RESB=1
for ring in 3 4 5; do
timeout 300 python tools/gemm_quick.py --tag cw64_ring$ring --ring $ring --which fc1,dz,fc2,dxhat --stages 2 2>&1 | tail -1
done | tee gpurun_out/ring.log
for ring in 3 4 5 6; do
CNX_LIB_PATH=$PWD/gpurun_in/libcnx_cw32.so timeout 300 python tools/gemm_quick.py --tag cw32_ring$ring --ring $ring --which fc1,dz,fc2,dxhat --stages 2 2>&1 | tail -1
done | tee -a gpurun_out/ring.log
