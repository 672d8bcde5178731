#!/usr/bin/env bash
# V4 smoke test (ref final_project/v4_mpi_cuda/test_v4.sh:15-16,82-123): build, run np in NP_VALUES,
# pass = exit 0 AND no "WARNING:" line in the log; prints the times.  Logs go to logs_v4_test/.
set -uo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ROOT="$(cd "$HERE/../.." && pwd)"
NP_VALUES=(${NP_VALUES:-1 2 4})
MPI_ARGS=${MPI_ARGS:---oversubscribe}
LOGDIR="$HERE/logs_v4_test"
mkdir -p "$LOGDIR"
export PATH="$ROOT/bin:$PATH"
cd "$HERE"
if ! make clean >/dev/null 2>&1 || ! make > "$LOGDIR/build.log" 2>&1; then
  echo "[FAIL] build (see $LOGDIR/build.log)"; exit 1
fi
echo "[ OK ] build"
fail=0
for np in "${NP_VALUES[@]}"; do
  log="$LOGDIR/v4_np${np}.log"
  if mpirun $MPI_ARGS -np "$np" ./template "$@" > "$log" 2>&1; then
    if grep -q "WARNING:" "$log"; then
      echo "[WARN] np=$np produced warnings (see $log)"; fail=1
    else
      t=$(grep -oE "completed in [0-9.]+ ms" "$log" | head -1)
      echo "[ OK ] np=$np  $t  $(grep -m1 '^Final Output Shape:' "$log")"
    fi
  else
    echo "[FAIL] np=$np exit code $? (see $log)"; fail=1
  fi
done
exit $fail
