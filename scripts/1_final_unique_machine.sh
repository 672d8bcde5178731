#!/usr/bin/env bash
# Full local suite: V1; V2.1 & V2.2 x np{1,2,4}; V3; V4 (and V5) x np{1,2,4}
# (ref scripts/1_final_unique_machine.sh).  Thin wrapper: the build-run-parse engine is b200cnn/harness/suite.py.
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
export PYTHONPATH="$ROOT${PYTHONPATH:+:$PYTHONPATH}"
exec python -m b200cnn.harness.suite --session-prefix script1 "$@"
