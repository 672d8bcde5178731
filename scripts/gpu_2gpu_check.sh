#!/usr/bin/env bash
# Two GPUs: the multi-process tests and the bench line (gpurun --gpus 2 -- 'bash scripts/gpu_2gpu_check.sh').
set -u
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"; cd "$ROOT"
OUT=gpurun_out/np2; mkdir -p $OUT
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:${LD_LIBRARY_PATH:-}
T0=$(date +%s)
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29602 bench.py --gpus 2 --steps 20 --warmup 5 --alt 2>$OUT/bench_2.err | grep "^{" | tee $OUT/bench_ours_2.json | cut -c1-200
echo "[$(( $(date +%s) - T0 )) s]"
timeout 400 python -m pytest tests/test_multigpu.py -x -q 2>&1 | tee $OUT/pytest_multigpu.log | tail -n 5
echo "[$(( $(date +%s) - T0 )) s]"
