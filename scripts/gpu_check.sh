#!/usr/bin/env bash
# One-shot GPU validation for a gpurun call: tests, smoke, short bench.  Never aborts early:
# every stage logs to gpurun_out/ so a single call returns as much evidence as possible.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
if [ -n "${TC_DEBUG:-}" ]; then
  echo "== tc_debug"; timeout 300 python scripts/tc_debug.py 2>&1 | tee gpurun_out/tc_debug.log | tail -60
fi
echo "== pytest -m gpu"; timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tee gpurun_out/pytest_gpu.log | tail -25
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tee gpurun_out/smoke.log | tail -8
echo "== bench"; timeout 600 python bench.py --steps "${STEPS:-10}" --warmup 3 ${BENCH_ARGS:-} 2>gpurun_out/bench.err | tee gpurun_out/bench.json
tail -5 gpurun_out/bench.err
