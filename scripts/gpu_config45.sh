#!/usr/bin/env bash
# BASELINE.json configs 4 and 5 on one 8-GPU box: 2048x2048x3 batch 16 (halo-dominated strong scaling)
# stage timings, and the full network (spatial blocks 1-2 + tensor-parallel FC6-8) at batch 256.
set -u
cd "$(dirname "${BASH_SOURCE[0]}")/.."
OUT=gpurun_out/config45; mkdir -p $OUT
NG=$(nvidia-smi -L | wc -l)
for n in 1 $NG; do
  HW=2048 BATCH=16 ROWGROUPS=rows HALOS=push,pull STRATEGIES=halo,recompute timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2981$n scripts/mgpu_experiments.py 2>$OUT/c4_$n.err | tee $OUT/config4_2048_np$n.log | grep -E "^---|rank 0:|slowest" | cut -c1-170
done
for n in 1 $NG; do
  PYTHONPATH=. timeout 400 bin/mpirun -np $n python -m b200cnn.runners.full --batch 256 --iters 10 --row-groups auto 2>$OUT/c5_$n.err | tee $OUT/config5_full_np$n.log | grep -E "completed|device_forward|images_per_s|argmax0"
done
