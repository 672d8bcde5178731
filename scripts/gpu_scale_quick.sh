#!/usr/bin/env bash
# bench.py at N = 2, 4, 8 on one 8-GPU box (N=1 comes from the 1-GPU runs); results in gpurun_out/scale2/
set -u
cd "$(dirname "${BASH_SOURCE[0]}")/.."
OUT=gpurun_out/scale2; mkdir -p $OUT
for n in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2970$n bench.py --gpus $n --steps 20 --warmup 5 2>$OUT/bench_$n.err | tee $OUT/bench_ours_$n.json | cut -c1-200
done
