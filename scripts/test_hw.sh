#!/usr/bin/env bash
# test_hw.sh <N> — homework toolchain (ref scripts/test_hw.sh); engine: b200cnn/harness/hw.py
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
export PYTHONPATH="$ROOT${PYTHONPATH:+:$PYTHONPATH}"
exec python -m b200cnn.harness.hw test "$@"
