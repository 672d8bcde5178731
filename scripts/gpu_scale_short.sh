#!/usr/bin/env bash
# The few multi-GPU numbers that matter most, in under five minutes of box time
# (gpurun --gpus 8 --timeout 300 -- 'bash scripts/gpu_scale_short.sh').  Logs under gpurun_out/r2_8gpu/.
set -u
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"; cd "$ROOT"
OUT=gpurun_out/r2_8gpu; mkdir -p $OUT
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:${LD_LIBRARY_PATH:-}
NG=$(nvidia-smi -L | wc -l)
T0=$(date +%s)
trun() { local t=$1 n=$2 port=$3; shift 3; timeout $t python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@"; }
stamp() { echo "[$(( $(date +%s) - T0 )) s] $*"; }
for n in 8 4; do
  [ $n -le $NG ] || continue
  stamp "bench ours N=$n"
  trun 110 $n 2960$n bench.py --gpus $n --steps 20 --warmup 5 --alt 2>$OUT/bench_$n.err | grep "^{" | tee $OUT/bench_ours_$n.json | cut -c1-200
done
[ 8 -le $NG ] && { stamp "stage timings N=8"; HALOS=push STRATEGIES=halo trun 60 8 29618 scripts/mgpu_experiments.py 2>$OUT/mg_8.err | tee $OUT/mgpu_experiments_np8.log | grep -E "^---|slowest" | cut -c1-170; }
[ 8 -le $NG ] && { stamp "bench reference N=8"; trun 60 8 29628 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 2>$OUT/benchref_8.err | grep "^{" | tee $OUT/bench_reference_8.json | cut -c1-300; }
stamp "bench ours N=2"
trun 80 2 29602 bench.py --gpus 2 --steps 20 --warmup 5 --alt 2>$OUT/bench_2.err | grep "^{" | tee $OUT/bench_ours_2.json | cut -c1-200
stamp done
