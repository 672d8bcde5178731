#!/usr/bin/env bash
# Dump every source file into project.txt (ref collect_project.sh -> project.txt) for review tools.
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"; OUT="${1:-$ROOT/project.txt}"
: > "$OUT"
cd "$ROOT"
git ls-files | grep -E '\.(py|cu|cuh|cpp|c|h|sh|mk|md|template)$|Makefile$|bin/mpirun$' | grep -v '^SURVEY.md$' | while read -r f; do
  printf '\n===== %s =====\n' "$f" >> "$OUT"; cat "$f" >> "$OUT"
done
echo "wrote $OUT ($(wc -l < "$OUT") lines)"
