#!/usr/bin/env bash
# Everything that needs the 8-GPU box, most important first (run with: gpurun --gpus 8 --timeout 1500 -- 'bash scripts/gpu_round2_all.sh').
# Logs under gpurun_out/r2_8gpu/.
set -u
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"; cd "$ROOT"
OUT=gpurun_out/r2_8gpu; mkdir -p $OUT
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:${LD_LIBRARY_PATH:-}
NG=$(nvidia-smi -L | wc -l)
run() { timeout ${TMO:-300} python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
nvidia-smi topo -m > $OUT/topo.txt 2>&1
for n in 8 4 2 1; do
  [ $n -le $NG ] || continue
  echo "== bench ours N=$n"
  if [ $n -eq 1 ]; then timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 2>$OUT/bench_$n.err | tee $OUT/bench_ours_$n.json | cut -c1-260
  else TMO=300 run $n 2960$n bench.py --gpus $n --steps 20 --warmup 5 --alt 2>$OUT/bench_$n.err | grep "^{" | tee $OUT/bench_ours_$n.json | cut -c1-260; fi
done
for n in 1 2 4 8; do
  [ $n -le $NG ] || continue
  echo "== bench reference N=$n"
  if [ $n -eq 1 ]; then timeout 300 python bench.py --impl reference --gpus 1 --steps 5 --warmup 1 2>$OUT/benchref_$n.err | tee $OUT/bench_reference_$n.json | cut -c1-200
  else TMO=300 run $n 2962$n bench.py --impl reference --gpus $n --steps 5 --warmup 1 2>$OUT/benchref_$n.err | grep "^{" | tee $OUT/bench_reference_$n.json | cut -c1-200; fi
done
for n in 8 4; do
  [ $n -le $NG ] || continue
  echo "== stage timings N=$n"
  HALOS=push,pull STRATEGIES=halo,recompute TMO=300 run $n 2961$n scripts/mgpu_experiments.py 2>$OUT/mg_$n.err | tee $OUT/mgpu_experiments_np$n.log | grep -E "^---|rank 0|rank 1:|slowest" | cut -c1-170
done
echo "== multi-GPU tests (np = all GPUs)"; timeout 600 python -m pytest tests/test_multigpu.py -x -q -k "all_gpus" 2>&1 | tee $OUT/pytest_multigpu.log | tail -3
echo "== config 4 (2048x2048x3, batch 16) and 5 (full network, batch 256)"
for n in 1 $NG; do
  if [ $n -eq 1 ]; then HW=2048 BATCH=16 timeout 200 python scripts/mgpu_experiments.py 2>$OUT/c4_$n.err | tee $OUT/config4_2048_np$n.log | grep -E "^---|slowest" | cut -c1-170
  else HW=2048 BATCH=16 ROWGROUPS=rows HALOS=push,pull STRATEGIES=halo,recompute TMO=300 run $n 2981$n scripts/mgpu_experiments.py 2>$OUT/c4_$n.err | tee $OUT/config4_2048_np$n.log | grep -E "^---|slowest" | cut -c1-170; fi
done
for n in 1 $NG; do
  PYTHONPATH=. timeout 300 bin/mpirun -np $n python -m b200cnn.runners.full --batch 256 --iters 10 --row-groups auto 2>$OUT/c5_$n.err | tee $OUT/config5_full_np$n.log | grep -E "completed|device_forward|images_per_s|argmax0"
done
echo "== HW1 n=4096 (FP64 GEMM rows over peer memory)"
for n in 1 2 4 8; do
  [ $n -le $NG ] || continue
  PYTHONPATH=. timeout 200 bin/mpirun -np $n python -m b200cnn.runners.hw1 4096 --backend gpu 2>$OUT/hw1_$n.err | grep -E "Parallel computation|Test:" | sed "s/^/np=$n /" | tee -a $OUT/hw1_n4096.log
done
echo "== same-box baselines"
OUTB=$OUT bash scripts/gpu_baselines.sh 2>&1 | tail -60
