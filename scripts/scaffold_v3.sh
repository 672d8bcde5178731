#!/usr/bin/env bash
# Clone a version directory into a new one and rename its launcher target
# (ref scripts/scaffold_v3.sh: clones V1 into V3 and renames files).  Usage: scaffold_v3.sh <src> <dst> <runner>
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
SRC="${1:-v3_cuda_only}"; DST="${2:?destination directory under final_project/}"; RUNNER="${3:-b200cnn.runners.v3}"
[ -e "$ROOT/final_project/$DST" ] && { echo "final_project/$DST exists"; exit 1; }
mkdir -p "$ROOT/final_project/$DST"
sed "s|^RUNNER := .*|RUNNER := $RUNNER|" "$ROOT/final_project/$SRC/Makefile" > "$ROOT/final_project/$DST/Makefile"
echo "scaffolded final_project/$DST (runner $RUNNER)"
