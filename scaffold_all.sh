#!/usr/bin/env bash
# Re-creates the repository skeleton (ref scaffold_all.sh): directories, empty package markers and
# per-version Makefiles.  Idempotent; never overwrites an existing file.
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
P="$ROOT/cuda-mpi-gpu-cluster-programming_b200"
for d in "$P"/{models,ops,parallel,utils,runners,harness} "$ROOT"/{b200cnn,csrc/cpu,csrc/cuda,csrc/apps,include,tests,scripts,bin,templates,baseline,profiles,homeworks} \
         "$ROOT"/final_project/{v1_serial,v2_mpi_only/2.1_broadcast_all,v2_mpi_only/2.2_scatter_halo,v3_cuda_only,v4_mpi_cuda,v5_cuda_aware_mpi}; do
  mkdir -p "$d"
done
for d in models ops parallel utils runners harness; do [ -e "$P/$d/__init__.py" ] || : > "$P/$d/__init__.py"; done
echo "skeleton present under $ROOT"
