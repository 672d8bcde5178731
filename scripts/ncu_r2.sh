#!/bin/bash
# ncu --set full of the block kernels the bench runs: one GPU (WORLD=1) and the launch-bound share of one rank of an
# 8-GPU run (WORLD=8), each after a plain run of the same command.   gpurun --timeout 900 -- 'bash scripts/ncu_r2.sh'
out=gpurun_out/ncu_r2; mkdir -p $out
for W in 1 8; do
  skip=$( [ $W -eq 1 ] && echo 2 || echo 16 )
  WORLD=$W timeout 200 python scripts/ncu_rank_case.py > $out/plain_w$W.log 2>&1 || { echo "plain run W=$W failed"; tail -5 $out/plain_w$W.log; exit 1; }
  WORLD=$W timeout 400 ncu --set full --clock-control none --import-source on -k regex:'tc_block' -s $skip -c 2 \
      -f -o $out/prof_w$W python scripts/ncu_rank_case.py > $out/ncu_w$W.log 2>&1
  echo "ncu W=$W exit $?"; tail -n 2 $out/ncu_w$W.log
done
