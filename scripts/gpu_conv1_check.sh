#!/usr/bin/env bash
# One GPU: conv1 with all-TMA operands + register pooling against the previous loader / epilogue.
set -u
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"; cd "$ROOT"
OUT=gpurun_out/conv1; mkdir -p $OUT
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:${LD_LIBRARY_PATH:-}
echo "== variants (0 = shipped, 1024 = cp.async tails, 2048 = strips, 3072 = both = previous kernel)"
B200_TC_VERBOSE=1 KNOBS=0,1024,2048,3072,0 timeout 120 python scripts/conv1_knobs.py 2>&1 | tee $OUT/variants.log | sort | uniq -c | sort -k2 | head -30
echo "== decomposition of the shipped kernel"
timeout 120 python scripts/conv1_knobs.py 2>&1 | tee $OUT/knobs.log
echo "== tests"
timeout 400 python -m pytest tests/test_zy_db_kernel.py tests/test_gpu_tc.py -x -q 2>&1 | tee $OUT/pytest.log | tail -n 12
echo "== bench"
timeout 150 python bench.py --steps 30 --warmup 5 2>$OUT/bench.err | tee $OUT/bench.json | cut -c1-230
