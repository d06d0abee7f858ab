#!/bin/bash
# A/B of an environment switch on the full bench in ONE box: gpu_ab.sh VAR  (runs tests first)
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
VAR="${1:-CNX_NO_PDL}"
timeout 1200 python -m pytest tests -q -m gpu -p no:cacheprovider -x > gpurun_out/test_all.log 2>&1
echo "== gpu tests exit=$? : $(tail -1 gpurun_out/test_all.log)"
grep -E "^(E  |FAILED)" gpurun_out/test_all.log | cut -c1-300 | head -20
for i in 1 2; do
  timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_a$i.log 2> gpurun_out/bench_a$i.err
  echo "== default exit=$? $(python -c "import json;d=json.loads(open('gpurun_out/bench_a$i.log').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['e2e']['value'])")"
  env $VAR=1 timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_b$i.log 2> gpurun_out/bench_b$i.err
  echo "== $VAR=1 exit=$? $(python -c "import json;d=json.loads(open('gpurun_out/bench_b$i.log').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['e2e']['value'])")"
done
