#!/usr/bin/env bash
# Sourceable harness library: the function names scripts written against the reference's
# scripts/common_test_utils.sh call (ref scripts/common_test_utils.sh:13-346), backed by the Python
# harness (b200cnn/harness/suite.py) instead of bash string handling.
#
#   source scripts/common_test_utils.sh
#   init_test_session mysuite                      # SESSION_ID, SESSION_LOG_DIR, CSV_SUMMARY_FILE, ...
#   _detect_and_set_cuda_arch_flags                # exports DETECTED_GPU_ARCH_FLAGS (sm_100a on B200)
#   setup_test_case "V4 MPI+CUDA" 2 v4_mpi_cuda v4_mpi_cuda && execute_single_test_case
#   _print_summary_table
#
# Row format, log names (make_<tag>_npN.log / run_<tag>_npN.log), pass/fail classification and the
# 20-column CSV are those of the suite scripts 0/1/2.
_B200_ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
export PYTHONPATH="$_B200_ROOT${PYTHONPATH:+:$PYTHONPATH}"
FP="${FP:-$_B200_ROOT/final_project}"
LOGS_BASE_DIR="${LOGS_BASE_DIR:-$FP/logs}"

init_test_session() {                              # arg: session prefix (default "custom")
    local ts; ts=$(date +%Y%m%d_%H%M%S)
    MACHINE_ID_CLEAN=$(hostname | tr -c 'A-Za-z0-9_.-' '_')
    GIT_COMMIT_SHORT=$(git -C "$_B200_ROOT" rev-parse --short HEAD 2>/dev/null || echo unknown)
    SESSION_ID="${1:-custom}_session_${ts}_${MACHINE_ID_CLEAN}"
    SESSION_LOG_DIR="$LOGS_BASE_DIR/$SESSION_ID"
    CSV_SUMMARY_FILE="$SESSION_LOG_DIR/summary_report_${ts}.csv"
    mkdir -p "$SESSION_LOG_DIR"
    echo "[harness] session $SESSION_ID -> $SESSION_LOG_DIR"
}

_detect_and_set_cuda_arch_flags() {
    DETECTED_GPU_ARCH_FLAGS="$(python -m b200cnn.harness.suite arch)"
    export DETECTED_GPU_ARCH_FLAGS
    echo "  [ARCH_DETECT_FINAL] DETECTED_GPU_ARCH_FLAGS globally set to: '$DETECTED_GPU_ARCH_FLAGS'"
}

setup_test_case() {                                # 1: display name, 2: np, 3: dir under final_project, 4: file tag
    CURRENT_VARIANT_NAME="$1"; CURRENT_NP="$2"; CURRENT_VARIANT_SUBDIR="$3"; CURRENT_VARIANT_TAG="$4"
    PROJECT_VARIANT_DIR="$FP/$3"
    echo "=== ${CURRENT_VARIANT_NAME} (NP=${CURRENT_NP}) in ${PROJECT_VARIANT_DIR} [${CURRENT_VARIANT_TAG}]"
    [ -n "${SESSION_LOG_DIR:-}" ] || init_test_session custom
    if [ ! -d "$PROJECT_VARIANT_DIR" ]; then           # logged as a failed row, like the reference does
        execute_single_test_case
        return 1
    fi
}

execute_single_test_case() {                       # extra args are passed to ./template
    if [ -z "${CURRENT_VARIANT_NAME:-}" ] || [ -z "${SESSION_LOG_DIR:-}" ]; then
        echo "[ERROR] Test case not properly set up. Call setup_test_case first."; return 1
    fi
    python -m b200cnn.harness.suite case --name "$CURRENT_VARIANT_NAME" --np "$CURRENT_NP" \
        --subdir "$CURRENT_VARIANT_SUBDIR" --tag "$CURRENT_VARIANT_TAG" \
        --session-dir "$SESSION_LOG_DIR" --csv "$CSV_SUMMARY_FILE" --timeout "${TEST_TIMEOUT:-600}" -- "$@"
}

_log_to_csv() {                                    # 16 positional fields after the 4 session columns
    local ts; ts=$(date --iso-8601=seconds)
    [ -s "$CSV_SUMMARY_FILE" ] || python -c "from b200cnn.harness.suite import CSV_HEADER; print(','.join(CSV_HEADER))" > "$CSV_SUMMARY_FILE"
    python - "$CSV_SUMMARY_FILE" "$SESSION_ID" "$MACHINE_ID_CLEAN" "$GIT_COMMIT_SHORT" "$ts" "$@" <<'PY'
import csv, sys
with open(sys.argv[1], "a", newline="") as f:
    csv.writer(f).writerow(sys.argv[2:22])
PY
}

_print_summary_table() {
    echo ""; echo "=== Summary Table (Machine: ${MACHINE_ID_CLEAN:-?}, Session: ${SESSION_ID:-?}) ==="
    python -m b200cnn.harness.suite table "$CSV_SUMMARY_FILE"
}
