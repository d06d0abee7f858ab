#!/bin/bash
# 2-GPU sweep of the gradient-exchange knobs (bucket size, NCCL CTA count, graphs on/off).
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
run() {   # tag, env...
  tag=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29620 bench.py --gpus 2 --steps 10 --warmup 3 --no-configs --no-e2e --no-cpu-baseline > gpurun_out/sweep_$tag.log 2> gpurun_out/sweep_$tag.err
  python - "$tag" <<'PY'
import json, sys
tag = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/sweep_{tag}.log").read().strip().splitlines()[-1])
    print(f"{tag:28s} {d['value']:9.1f} img/s  {d['ms_per_step']:7.3f} ms  identical={d.get('replicas_identical')}")
except Exception as e:
    print(tag, "FAILED", e)
PY
}
python bench.py --steps 10 --warmup 3 --no-configs --no-e2e --no-cpu-baseline > gpurun_out/sweep_1gpu.log 2>/dev/null
python -c "import json; d=json.loads(open('gpurun_out/sweep_1gpu.log').read().strip().splitlines()[-1]); print('1 GPU                        %9.1f img/s  %7.3f ms' % (d['value'], d['ms_per_step']))"
run default CNX_BUCKET_MB=32
run bucket128 CNX_BUCKET_MB=128
run bucket512 CNX_BUCKET_MB=512
run ctas4 CNX_BUCKET_MB=32 NCCL_MAX_CTAS=4
run ctas2_b128 CNX_BUCKET_MB=128 NCCL_MAX_CTAS=2
run nograph CNX_BUCKET_MB=32 CNX_NO_GRAPH=1
