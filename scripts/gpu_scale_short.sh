#!/usr/bin/env bash
# The few multi-GPU numbers that matter most, in about two minutes of box time
# (gpurun --gpus 8 --timeout 200 -- 'bash scripts/gpu_scale_short.sh').  Logs under gpurun_out/r2_8gpu_v2/.
set -u
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"; cd "$ROOT"
OUT=gpurun_out/r2_8gpu_v2; mkdir -p $OUT
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:${LD_LIBRARY_PATH:-}
NG=$(nvidia-smi -L | wc -l)
T0=$(date +%s)
trun() { local t=$1 n=$2 port=$3; shift 3; timeout $t python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@"; }
stamp() { echo "[$(( $(date +%s) - T0 )) s] $*"; }
for n in 8 4; do
  [ $n -le $NG ] || continue
  stamp "bench ours N=$n"
  trun 90 $n 2960$n bench.py --gpus $n --steps 20 --warmup 5 --alt 2>$OUT/bench_$n.err | grep "^{" | tee $OUT/bench_ours_$n.json | cut -c1-200
done
[ 8 -le $NG ] && { stamp "stage timings N=8"; HALOS=push STRATEGIES=halo,recompute trun 60 8 29618 scripts/mgpu_experiments.py 2>$OUT/mg_8.err | tee $OUT/mgpu_experiments_np8.log | grep -E "^---|slowest" | cut -c1-170; }
for n in 8 4 2; do
  [ $n -le $NG ] || continue
  stamp "host link ceiling N=$n"
  trun 40 $n 2972$n scripts/h2d_roofline.py 2>$OUT/h2d_$n.err | tee $OUT/h2d_roofline_np$n.json | cut -c1-300
done
[ 8 -le $NG ] && { stamp "config 4 (2048x2048x3, batch 16) N=8"; HW=2048 BATCH=16 ROWGROUPS=rows HALOS=push,pull STRATEGIES=halo,recompute trun 80 8 29818 scripts/mgpu_experiments.py 2>$OUT/c4_8.err | tee $OUT/config4_2048_np8.log | grep -E "^---|slowest" | cut -c1-170; }
stamp "config 5 (full network, batch 256) N=$NG"
PYTHONPATH=. timeout 90 bin/mpirun -np $NG python -m b200cnn.runners.full --batch 256 --iters 10 --row-groups auto 2>$OUT/c5_$NG.err | tee $OUT/config5_full_np$NG.log | grep -E "completed|device_forward|images_per_s"
stamp done
