#!/bin/bash
# Round-end pass on ONE GPU: whole -m gpu suite, smoke, bench (224 and 384), reference arm.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/test_all.log 2>&1
echo "== gpu tests exit=$? : $(tail -1 gpurun_out/test_all.log)"
grep -E "^(E  |FAILED)" gpurun_out/test_all.log | cut -c1-300 | head -20
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "== smoke exit=$?"
timeout 900 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err
echo "== bench exit=$?"; tail -3 gpurun_out/bench.err; cut -c1-900 gpurun_out/bench.log
timeout 900 python bench.py --res 384 --micro-batch 64 --no-cpu-baseline > gpurun_out/bench384.log 2> gpurun_out/bench384.err
echo "== bench384 exit=$?"; tail -3 gpurun_out/bench384.err; cut -c1-500 gpurun_out/bench384.log
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err
echo "== reference arm exit=$?"; tail -3 gpurun_out/bench_ref.err; cut -c1-700 gpurun_out/bench_ref.log
