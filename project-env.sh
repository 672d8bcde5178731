#!/usr/bin/env bash
# Environment check + full run (ref shell.nix / project-nix-shell.sh:9-13, which pin a Nix shell and
# then run scripts 0, 1, 2).  This image has no Nix: the toolchain is verified instead.
set -uo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ok=1
for t in python g++ make nvcc; do
  if command -v $t >/dev/null; then printf '%-8s %s\n' "$t" "$($t --version 2>&1 | head -1)"; else echo "$t MISSING"; ok=0; fi
done
python - <<'PY' || ok=0
import torch, pandas
print("torch", torch.__version__, "cuda", torch.version.cuda, "gpus", torch.cuda.device_count())
PY
[ $ok -eq 1 ] || { echo "environment incomplete"; exit 2; }
[ "${1:-}" = "--check" ] && exit 0
export PATH="$ROOT/bin:$PATH"
"$ROOT/scripts/0_run_final_project.sh" && "$ROOT/scripts/1_final_unique_machine.sh" && "$ROOT/scripts/2_final_multi_machine.sh" --dry-run
