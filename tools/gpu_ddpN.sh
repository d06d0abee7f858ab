#!/bin/bash
# N-GPU bench (N = first argument), plus ddp_equiv at N ranks.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
N=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${N}gpu.log 2> gpurun_out/bench_${N}gpu.err
echo "== bench ${N}gpu exit=$?"; tail -3 gpurun_out/bench_${N}gpu.err | cut -c1-300; cut -c1-300 gpurun_out/bench_${N}gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 tools/ddp_equiv.py > gpurun_out/ddp_equiv_${N}.log 2> gpurun_out/ddp_equiv_${N}.err
echo "== ddp_equiv exit=$?"; tail -2 gpurun_out/ddp_equiv_${N}.err | cut -c1-300; tail -1 gpurun_out/ddp_equiv_${N}.log
