#!/usr/bin/env bash
# Concatenate the documentation into one file (ref collect_p_docs.sh builds the root README).
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"; OUT="${1:-$ROOT/docs_bundle.md}"
: > "$OUT"
for f in README.md final_project/README.md final_project/docs/AlexNet_B200_Design.md DESIGN.md INTEGRATION.md \
         COMPONENTS.md BASELINE.md profiles/README.md profiles/ROOFLINE.md homeworks/hw1/summary.md; do
  [ -f "$ROOT/$f" ] && { printf '\n\n<!-- %s -->\n\n' "$f" >> "$OUT"; cat "$ROOT/$f" >> "$OUT"; }
done
echo "wrote $OUT"
