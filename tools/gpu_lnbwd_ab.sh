#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_block.py tests/test_gpu_fullsize.py tests/test_gpu_model.py tests/test_gpu_loop.py -q -m gpu -x -p no:cacheprovider > gpurun_out/test_ln.log 2>&1
echo "== tests exit=$? : $(tail -1 gpurun_out/test_ln.log)"
grep -E "^(E  |FAILED)" gpurun_out/test_ln.log | cut -c1-400 | head
for v in 0 1 0 1; do
  CNX_LNBWD_ROWKERNEL=$v python bench.py --steps 10 --warmup 3 --no-configs --no-e2e --no-cpu-baseline > gpurun_out/ln_$v.log 2> gpurun_out/ln_$v.err
  python -c "import json; d=json.loads(open('gpurun_out/ln_$v.log').read().strip().splitlines()[-1]); print('rowkernel=$v  %9.1f img/s  %7.3f ms  launches %d' % (d['value'], d['ms_per_step'], d['gpu_launches']))" || tail -3 gpurun_out/ln_$v.err
done
