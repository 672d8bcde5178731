#!/usr/bin/env bash
# Multi-GPU validation + scaling on one box (run with gpurun --gpus 8): multigpu tests, bench at
# N = 1,2,4,8 (ours, then the reference arm), stage timings at every N, NCCL/V5/ref-V4 baselines.
set -u
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"; cd "$ROOT"
OUT=gpurun_out/scale; mkdir -p $OUT
NG=$(nvidia-smi -L | wc -l)
nvidia-smi topo -m > $OUT/topo.txt 2>&1
echo "== multigpu tests"; timeout 900 python -m pytest tests/test_multigpu.py -x -q 2>&1 | tee $OUT/pytest_multigpu.log | tail -4
for n in 1 2 4 8; do
  [ $n -le $NG ] || continue
  echo "== bench ours N=$n"
  if [ $n -eq 1 ]; then timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 2>$OUT/bench_$n.err | tee $OUT/bench_ours_$n.json
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2960$n bench.py --gpus $n --steps 20 --warmup 5 --alt 2>$OUT/bench_$n.err | tee $OUT/bench_ours_$n.json; fi
  echo "== stage timings N=$n"
  HALOS=push,pull timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n scripts/mgpu_experiments.py 2>$OUT/mg_$n.err | tee $OUT/mgpu_experiments_np$n.log | grep -E "^---|rank 0|rank 1:|slowest" | cut -c1-170
done
SKIP_V1=1 bash scripts/gpu_baselines.sh 2>&1 | tail -40
for n in 1 2 4 8; do
  [ $n -le $NG ] || continue
  echo "== bench reference N=$n"
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2962$n bench.py --impl reference --gpus $n --steps 5 --warmup 1 2>$OUT/benchref_$n.err | tee $OUT/bench_reference_$n.json | cut -c1-200
done
