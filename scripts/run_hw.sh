#!/usr/bin/env bash
# run_hw.sh <N> — homework toolchain (ref scripts/run_hw.sh); engine: b200cnn/harness/hw.py
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
export PYTHONPATH="$ROOT${PYTHONPATH:+:$PYTHONPATH}"
exec python -m b200cnn.harness.hw run "$@"
