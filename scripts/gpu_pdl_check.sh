#!/usr/bin/env bash
# One GPU: block 2 as a programmatic dependent launch of block 1, off vs on (bench line + bit-identity tests).
set -u
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"; cd "$ROOT"
OUT=gpurun_out/pdl; mkdir -p $OUT
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:${LD_LIBRARY_PATH:-}
for v in 0 1 0 1; do
  echo "== bench B200_TC_PDL=$v"
  B200_TC_PDL=$v timeout 150 python bench.py --steps 30 --warmup 5 2>$OUT/bench_pdl$v.err | tee -a $OUT/bench_pdl$v.json | cut -c1-230
done
echo "== tests with the default policy"
timeout 400 python -m pytest tests/test_zy_db_kernel.py tests/test_gpu_tc.py -x -q 2>&1 | tee $OUT/pytest.log | tail -n 5
echo "== host link ceiling, one GPU"
timeout 100 python scripts/h2d_roofline.py 2>$OUT/h2d.err | tee $OUT/h2d_roofline_np1.json
