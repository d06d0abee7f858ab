#!/bin/bash
# kernel timeline of one data-parallel optimizer step (torch.profiler): where the all-reduce sits
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
N=${1:-2}
for mb in 128 64; do
timeout 180 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29613 \
  tools/nccl_overlap.py --micro-batch $mb --out gpurun_out/nccl_overlap_${N}gpu_mb$mb.txt > gpurun_out/nccl_overlap_$mb.log 2>&1
echo "exit=$?"; tail -25 gpurun_out/nccl_overlap_$mb.log | cut -c1-160
done
