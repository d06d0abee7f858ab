#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_block.py tests/test_gpu_model.py tests/test_gpu_loop.py -q -m gpu -x -p no:cacheprovider > gpurun_out/test_side.log 2>&1
echo "== tests exit=$? : $(tail -1 gpurun_out/test_side.log)"
grep -E "^(E  |FAILED)" gpurun_out/test_side.log | cut -c1-300 | head
for v in 1 0 1; do
  CNX_BWD_SIDE_STREAM=$v python bench.py --steps 10 --warmup 3 --no-configs --no-e2e --no-cpu-baseline > gpurun_out/side_$v.log 2> gpurun_out/side_$v.err
  python -c "import json; d=json.loads(open('gpurun_out/side_$v.log').read().strip().splitlines()[-1]); print('side=$v  %9.1f img/s  %7.3f ms' % (d['value'], d['ms_per_step']))" || tail -3 gpurun_out/side_$v.err
done
