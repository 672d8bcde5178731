#!/usr/bin/env bash
# Full network on 2+ GPUs: conv3-5 on rank 0 vs sharded by batch (B200_FULL_CONV_SHARD), same probabilities,
# device-timed forward of each; plus an uneven batch (5 images) against the single-GPU run.
set -u
cd "$(dirname "${BASH_SOURCE[0]}")/.."
OUT=gpurun_out/full_shard; mkdir -p $OUT
NG=$(nvidia-smi -L | wc -l)
run() { tag=$1; np=$2; mode=$3; shift 3
  PYTHONPATH=. B200_FULL_CONV_SHARD=$mode timeout 200 bin/mpirun -np $np python -m b200cnn.runners.full "$@" > $OUT/$tag.log 2>$OUT/$tag.err
  echo "$tag rc $?: $(grep -E 'first 10 values' $OUT/$tag.log | cut -c1-120)"
  grep -oE "(device_forward_ms_warm|images_per_s_warm|argmax0|conv_shard)=[^ ]+" $OUT/$tag.log | tr '\n' ' '; echo; }
run b5_np1 1 rank0 --batch 5 --seed 3
run b5_np${NG}_batch $NG batch --batch 5 --seed 3
run b256_np${NG}_rank0 $NG rank0 --batch 256 --iters 10
run b256_np${NG}_batch $NG batch --batch 256 --iters 10
tail -3 $OUT/*.err | cut -c1-200 | tail -20
