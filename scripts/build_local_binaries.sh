#!/usr/bin/env bash
# Builds every variant and copies its `template` into prebuilt_executables_local/<variant>/
# (ref scripts/build_local_binaries.sh:47-116,162; the reference does this inside nix-shell).
set -uo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
OUT="$ROOT/prebuilt_executables_local"
variants=(v1_serial v2_mpi_only/2.1_broadcast_all v2_mpi_only/2.2_scatter_halo v3_cuda_only v4_mpi_cuda v5_cuda_aware_mpi)
rc=0
for v in "${variants[@]}"; do
  d="$ROOT/final_project/$v"
  echo "=== $v"
  if (cd "$d" && make clean >/dev/null && make); then
    mkdir -p "$OUT/$v" && cp "$d/template" "$OUT/$v/template" && echo "  -> $OUT/$v/template"
  else
    echo "  build FAILED"; rc=1
  fi
done
exit $rc
