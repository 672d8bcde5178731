#!/usr/bin/env bash
# Same-box baselines on a GPU box: reference V1 / V3 / V4 (unmodified sources,
# sm_100 rebuild), the bench's reference arm, and — with >= 2 GPUs — V5 reconstruction and the
# NCCL halo baseline.  Everything is logged under gpurun_out/baselines/.
set -u
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
cd "$ROOT"
OUT=${OUTB:-gpurun_out/baselines}; mkdir -p $OUT
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:${LD_LIBRARY_PATH:-}
NG=$(nvidia-smi -L | wc -l)
[ -x baseline/_ref/src/final_project/v4_mpi_cuda/template ] || python baseline/build_reference.py > $OUT/build.log 2>&1
R=baseline/_ref/src/final_project
[ -x baseline/v5_reconstruction/template ] || make -C baseline/v5_reconstruction > $OUT/v5_build.log 2>&1
echo "== ref V1 (CPU, 3 runs)"; for i in 1 2 3; do $R/v1_serial/template | grep "completed in"; done | tee $OUT/ref_v1.log
echo "== ref V3 (3 runs)"; for i in 1 2 3; do $R/v3_cuda_only/template | head -1; done | tee $OUT/ref_v3.log
echo "== reference forward functions in a warm loop (128 images after 3 warm-up calls; baseline/ref_loop)"
make -C baseline/ref_loop > $OUT/ref_loop_build.log 2>&1 || tail -3 $OUT/ref_loop_build.log
for v in v3 v4; do [ -x baseline/ref_loop/ref_loop_$v ] && baseline/ref_loop/ref_loop_$v 128 3; done | tee $OUT/ref_loop.log
for np in 1 2 4 8; do
  [ $np -le $NG ] || continue
  echo "== ref V4 np=$np (one GPU per rank)"
  for i in 1 2 3; do (cd $R/v4_mpi_cuda && B200_MPIRUN_BIND_GPUS=1 B200_MPIRUN_NUM_GPUS=$NG "$ROOT/bin/mpirun" --oversubscribe -np $np ./template 2>&1 | grep -E "Shape|completed"); done | tee $OUT/ref_v4_np$np.log
  if [ -x baseline/v5_reconstruction/template ]; then
    echo "== V5 reconstruction np=$np"
    for i in 1 2 3; do (cd baseline/v5_reconstruction && "$ROOT/bin/mpirun" --oversubscribe -np $np ./template 2>&1 | grep -E "Shape|completed"); done | tee $OUT/v5rec_np$np.log
  fi
  echo "== NCCL halo baseline np=$np"
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port 2954$np baseline/nccl_halo.py --check 2>$OUT/nccl_np$np.err | grep "^{" | tee $OUT/nccl_halo_np$np.json | cut -c1-300
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port 2955$np baseline/nccl_halo.py --check --sharded 2>>$OUT/nccl_np$np.err | grep "^{" | tee $OUT/nccl_halo_sharded_np$np.json | cut -c1-300
done

