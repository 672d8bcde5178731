#!/usr/bin/env bash
# Multi-machine suite (ref scripts/2_final_multi_machine.sh).  Configure with
#   B200_HOSTS="user@master:100a:8,user@worker:100a:8" [USE_SHARED_FILESYSTEM=true] [MPI_NETWORK_INTERFACE=eth0]
# Engine: b200cnn/harness/multi_machine.py (--dry-run prints the plan).
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
export PYTHONPATH="$ROOT${PYTHONPATH:+:$PYTHONPATH}"
exec python -m b200cnn.harness.multi_machine "$@"
