#!/bin/bash
# GEMM-extension check on one GPU: the full-AlexNet tests with the default tiling (N = 256 tiles + split-K), and,
# only if that fails, with each feature switched off, then the per-stage timings.
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 150 python -m pytest tests/test_gpu_full.py -x -q -m gpu > gpurun_out/gemm_tests_$tag.log 2>&1; rc=$?; echo "$tag rc $rc: $(tail -1 gpurun_out/gemm_tests_$tag.log)"; return $rc; }
if ! run default B200_X=0; then
    grep -m3 -n "Error\|error\|assert" gpurun_out/gemm_tests_default.log | cut -c1-200
    run bn128 B200_GEMM_BN128=1
    run nosplit B200_GEMM_SPLITK=0
    run bn128_nosplit B200_GEMM_BN128=1 B200_GEMM_SPLITK=0
fi
timeout 100 python scripts/full_stage_times.py 256 7 > gpurun_out/full_stage_times_v2.log 2>&1; echo "stage rc $?"
tail -12 gpurun_out/full_stage_times_v2.log
