#!/usr/bin/env bash
# Host-side memory / undefined-behaviour checking (AddressSanitizer + UBSan) of the native code that runs on the
# CPU: the oracle library and planner (csrc/cpu) through the native V1 and V2 binaries, and the MPI-subset shim (baseline/mpishim)
# through its self-test at np 1, 2 and 4 and the reference's HW1 under it.  compute-sanitizer is closed on the GPU
# pool, so this is the checked half; the device half relies on watchdogs and bit-exact multi-rank tests
# (DESIGN.md 4b).  The reference ships a .clang-tidy that nothing runs (ref final_project/v4_mpi_cuda/.clang-tidy:4).
#   scripts/sanitize_host.sh [outdir]      -> exit 0 when every run is clean; log in <outdir>/sanitize_host.log
set -u
cd "$(dirname "${BASH_SOURCE[0]}")/.."
OUT="${1:-/tmp/b200_sanitize_host}"; mkdir -p "$OUT"; LOG="$OUT/sanitize_host.log"; : > "$LOG"
SAN="-fsanitize=address,undefined -fno-sanitize-recover=undefined -fno-omit-frame-pointer -g -O1"
export ASAN_OPTIONS=detect_leaks=1:abort_on_error=0:exitcode=66 UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1
fail=0
run() { echo "\$ $*" >> "$LOG"; "$@" >> "$LOG" 2>&1; local rc=$?; echo "rc=$rc" >> "$LOG"; [ $rc -eq 0 ] || { echo "FAILED: $*"; fail=1; }; }

# 1. oracle library + native V1 entry point, warnings are errors
run g++ -std=c++17 $SAN -ffp-contract=off -fopenmp -Wall -Wextra -Werror -Iinclude \
    csrc/apps/v1_main.cpp csrc/cpu/*.cpp -o "$OUT/v1_asan"
run "$OUT/v1_asan" --seed 7 --quiet-extras
# ... and with parameters from a weight file through the native reader (csrc/cpu/weights_cpu.cpp)
run env PYTHONPATH=. python -c "from b200cnn import default_config; from b200cnn.config import INIT_RAND; from b200cnn.models.init import make_problem; from b200cnn.utils.checkpoint import save_weights; save_weights('$OUT/w.b2w', make_problem(default_config(init=INIT_RAND, seed=5))[1])"
run "$OUT/v1_asan" --seed 7 --quiet-extras --load-weights "$OUT/w.b2w"

# 1b. native V2.2 / V2.1 (planner, slab layers, scatter / halo / gather over the MPI subset) at np 1 / 3 / 4
run gcc $SAN -c -Wall -Wextra -Werror baseline/mpishim/mpishim.c -o "$OUT/mpishim_asan.o"
run g++ -std=c++17 $SAN -ffp-contract=off -fopenmp -Wall -Wextra -Werror -Iinclude -Ibaseline/mpishim \
    csrc/apps/v2_main.cpp csrc/cpu/*.cpp "$OUT/mpishim_asan.o" -lpthread -o "$OUT/v22_asan"
run g++ -std=c++17 $SAN -ffp-contract=off -fopenmp -Wall -Wextra -Werror -DV2_BROADCAST_ALL -Iinclude -Ibaseline/mpishim \
    csrc/apps/v2_main.cpp csrc/cpu/*.cpp "$OUT/mpishim_asan.o" -lpthread -o "$OUT/v21_asan"
for np in 1 3 4; do
  run python bin/mpirun --oversubscribe -np $np "$OUT/v22_asan" --height 99 --width 131 --batch 2 --init rand --seed 7 --quiet-extras
done
run python bin/mpirun --oversubscribe -np 2 "$OUT/v22_asan" --height 99 --width 99 --strategy recompute --quiet-extras
run python bin/mpirun --oversubscribe -np 2 "$OUT/v21_asan" --height 99 --width 99 --quiet-extras

# 2. MPI shim + its self-test, -Wall -Wextra -Werror, at np 1 / 2 / 4 through bin/mpirun
run gcc $SAN -Wall -Wextra -Werror -Ibaseline/mpishim tests/data/mpi_selftest.c baseline/mpishim/mpishim.c \
    -lpthread -o "$OUT/selftest_asan"
for np in 1 2 4; do
  run python bin/mpirun --oversubscribe -np $np "$OUT/selftest_asan"
done

# 3. the reference's HW1 (unmodified source) on the sanitized shim, n = 128, np 2
if [ -f /root/reference/homeworks/hw1/src/template.c ]; then
  run gcc $SAN -Ibaseline/mpishim /root/reference/homeworks/hw1/src/template.c baseline/mpishim/mpishim.c \
      -lpthread -lm -o "$OUT/hw1_asan"
  run python bin/mpirun --oversubscribe -np 2 "$OUT/hw1_asan" 128
fi
# 4. our own native HW1 (homeworks/hw1/src/template.c) on the sanitized shim and library sources, n = 128, np 1 / 4
run gcc -std=c11 $SAN -Wall -Wextra -Werror -Iinclude -Ibaseline/mpishim -c homeworks/hw1/src/template.c -o "$OUT/hw1n_asan.o"
run g++ -std=c++17 $SAN -fopenmp "$OUT/hw1n_asan.o" csrc/cpu/*.cpp "$OUT/mpishim_asan.o" -Iinclude -lpthread -lm -o "$OUT/hw1_native_asan"
for np in 1 4; do
  run python bin/mpirun --oversubscribe -np $np "$OUT/hw1_native_asan" 128
done
grep -c "ERROR: AddressSanitizer\|runtime error:" "$LOG" | sed 's/^/sanitizer reports: /'
[ $fail -eq 0 ] && echo "sanitize_host: clean" || echo "sanitize_host: FAILURES (see $LOG)"
exit $fail
