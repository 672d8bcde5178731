#!/bin/bash
# Per-stage device times (scripts/stage_times.py) of every conv1 / conv2 kernel variant, one process each,
# then ONE ncu --set full capture of the default kernels (CTA-pair conv2, pixel-TMA conv1).
# Run on the GPU box:  gpurun --timeout 600 -- 'bash scripts/variant_times.sh'
out=gpurun_out/r2b; mkdir -p $out
for v in "single_cpasync:B200_TC_CONV2=single B200_TC_CONV1=cpasync" "db_cpasync:B200_TC_CONV2=db B200_TC_CONV1=cpasync" "auto:"; do
  name=${v%%:*}; envs=${v#*:}
  echo "== $name ($envs)" | tee -a $out/stage_times.log
  env $envs timeout 120 python scripts/stage_times.py >> $out/stage_times.log 2>&1 || echo "FAILED $name rc=$?" >> $out/stage_times.log
done
cat $out/stage_times.log
timeout 200 python scripts/ncu_case.py > $out/ncu_case_plain.log 2>&1 || { echo "plain run failed"; tail -5 $out/ncu_case_plain.log; exit 1; }
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'tc_block' -s 2 -c 2 \
    -f -o $out/prof_pair_tma1 python scripts/ncu_case.py > $out/ncu.log 2>&1
echo "ncu exit $?"; tail -n 3 $out/ncu.log
