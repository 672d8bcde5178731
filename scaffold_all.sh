#!/usr/bin/env bash
# Creates the repository skeleton (ref scaffold_all.sh): directories, empty package markers and the
# per-version Makefiles.  Idempotent; never overwrites an existing file.
#   scaffold_all.sh            re-check this checkout
#   scaffold_all.sh <dir>      lay the skeleton down in <dir> (a new project that links against this one)
set -euo pipefail
SRC="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ROOT="${1:-$SRC}"
mkdir -p "$ROOT"; ROOT="$(cd "$ROOT" && pwd)"
P="$ROOT/cuda-mpi-gpu-cluster-programming_b200"
VERSIONS="v1_serial v2_mpi_only/2.1_broadcast_all v2_mpi_only/2.2_scatter_halo v3_cuda_only v4_mpi_cuda v5_cuda_aware_mpi"
for d in "$P"/{models,ops,parallel,utils,runners,harness} \
         "$ROOT"/{b200cnn,csrc/cpu,csrc/cuda,csrc/apps,include,tests,scripts,bin,templates,baseline,profiles,homeworks}; do
  mkdir -p "$d"
done
for d in models ops parallel utils runners harness; do [ -e "$P/$d/__init__.py" ] || : > "$P/$d/__init__.py"; done
for v in $VERSIONS; do
  mkdir -p "$ROOT/final_project/$v"
  # a new tree gets copies of the per-version Makefiles (they only name the runner module and include common.mk)
  if [ "$ROOT" != "$SRC" ] && [ ! -e "$ROOT/final_project/$v/Makefile" ] && [ -e "$SRC/final_project/$v/Makefile" ]; then
    cp "$SRC/final_project/$v/Makefile" "$ROOT/final_project/$v/Makefile"
  fi
done
if [ "$ROOT" != "$SRC" ]; then
  [ -e "$ROOT/final_project/common.mk" ] || cp "$SRC/final_project/common.mk" "$ROOT/final_project/common.mk"
  for t in "$SRC"/templates/*; do [ -e "$ROOT/templates/$(basename "$t")" ] || cp "$t" "$ROOT/templates/"; done
fi
echo "skeleton present under $ROOT"
