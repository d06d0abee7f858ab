#!/bin/bash
# 2-GPU pass: gradient / replica equivalence on hardware, then the 2-rank bench.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tools/ddp_equiv.py > gpurun_out/ddp_equiv.log 2> gpurun_out/ddp_equiv.err
echo "== ddp_equiv exit=$?"; tail -3 gpurun_out/ddp_equiv.err; tail -1 gpurun_out/ddp_equiv.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_2gpu.log 2> gpurun_out/bench_2gpu.err
echo "== bench 2gpu exit=$?"; tail -5 gpurun_out/bench_2gpu.err; cut -c1-600 gpurun_out/bench_2gpu.log
