#!/bin/bash
# ncu --set full of the final conv1 / conv2 / implicit-conv kernels (one ncu invocation per gpurun call).
# Run on the GPU box:  gpurun --timeout 400 -- 'bash scripts/ncu_final.sh'
mkdir -p gpurun_out
timeout 200 python scripts/ncu_case.py > gpurun_out/ncu_case_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_case_plain.log; exit 1; }
# launches matching the filter: [conv1, conv2] x 2 forwards, then the implicit conv x 2 -> skip the first forward
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'tc_block|tc_gemm' -s 2 -c 4 \
    -f -o gpurun_out/prof_final python scripts/ncu_case.py > gpurun_out/ncu_final.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_final.log
