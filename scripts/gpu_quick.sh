#!/usr/bin/env bash
# One GPU, under a minute: conv1 knob times, the kernel-equivalence tests, per-stage times and the bench line.
set -u
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"; cd "$ROOT"
OUT=gpurun_out/quick; mkdir -p $OUT
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:${LD_LIBRARY_PATH:-}
KNOBS=${KNOBS:-0,4096,4864,4544,5056} timeout 120 python scripts/conv1_knobs.py 2>&1 | tee $OUT/knobs.log
WORLDS=${WORLDS:-1,8} timeout 200 python scripts/rank_stage_times.py 2>&1 | tee $OUT/rank_stage_times.log
timeout 400 python -m pytest tests/test_zy_db_kernel.py tests/test_gpu_tc.py -x -q 2>&1 | tee $OUT/pytest.log | tail -n 6
timeout 150 python bench.py --steps 30 --warmup 5 2>$OUT/bench.err | tee $OUT/bench.json | cut -c1-230
