"""The log-line contract: exactly the stdout lines the reference binaries print
and its bash harness / log_analysis.py parse, plus compatible additions.

Rules the parsers impose (ref scripts/common_test_utils.sh:296-317):
* the time is the FIRST ``<num> ms`` anywhere in the log -> nothing before the canonical time line
  may contain that pattern; extra timings are printed as ``key_ms=<num>``;
* shape = ``^Final Output Shape: HxWxC`` else first ``Dimensions: H=, W=, C=`` else ``^shape: HxWxC``;
* first values = ``^Final Output (first 10 values):`` else ``^Sample values:``.
Numbers are formatted like C++ iostream defaults (``%g`` with 6 significant digits).
"""
from __future__ import annotations

import sys
from typing import Iterable, Optional, Sequence


def cxx_float(v: float) -> str:
    """std::cout << float with default precision(6)."""
    s = format(float(v), ".6g")
    if "e" in s:                      # iostream prints e-05, python prints e-05 too; normalise e+XX
        m, e = s.split("e")
        sign = e[0] if e[0] in "+-" else "+"
        digits = e.lstrip("+-").rjust(2, "0")
        s = f"{m}e{sign}{digits}"
    return s


def emit(line: str = "", file=None) -> None:
    try:
        print(line, file=file or sys.stdout, flush=True)
    except BrokenPipeError:           # `./template | head -3`: the reader is gone; keep running quietly
        import os
        try:
            sys.stdout = open(os.devnull, "w")
        except OSError:
            pass


# ---- V1 (ref v1_serial/src/main.cpp:14-75, alexnet_serial.cpp:61-185) ---------------------------
V1_BANNER = "--- AlexNet Serial CPU Version (V1) ---"
V1_LAYER_VERBS = {
    "After Conv1": "Applying Conv1...", "After ReLU1": "Applying ReLU1...",
    "After Pool1": "Applying MaxPool1...", "After Conv2": "Applying Conv2...",
    "After ReLU2": "Applying ReLU2...", "After Pool2": "Applying MaxPool2...",
    "After LRN2": "Applying LRN2...",
}


def v1_dimensions(layer: str, H: int, W: int, C: int) -> str:
    return f"  [{layer}] Dimensions: H={H}, W={W}, C={C}"


def v1_time(ms: float) -> str:
    return f"AlexNet Serial Forward Pass completed in {int(ms)} ms"          # integer ms


def v1_values(vals: Sequence[float], total: int) -> str:
    body = " ".join(cxx_float(v) for v in vals[:10])
    return f"Final Output (first 10 values): {body}" + ("..." if total > 10 else "")


# ---- V2.1 / V2.2 (ref v2_mpi_only/*/src/main.cpp:94-104, 269-285) --------------------------------
def v2_lines(H: int, W: int, C: int, vals: Sequence[float], ms: float):
    return [f"shape: {H}x{W}x{C}",
            "Sample values: " + " ".join(cxx_float(v) for v in vals[:5]),
            f"Execution Time: {cxx_float(ms)} ms"]


def v22_size_warning(got: int, H: int, W: int, C: int) -> str:
    return (f"WARNING: Gathered size ({got}) does not match expected final size ({H}x{W}x{C} = "
            f"{H * W * C}). Check trim/halo/layer logic.")


# ---- V3 (ref v3_cuda_only/src/main_cuda.cpp:36-41) ------------------------------------------------
def v3_lines(vals: Sequence[float], ms: float):
    return [f"AlexNet CUDA Forward Pass completed in {cxx_float(ms)} ms",
            "Final Output (first 10 values):" + "".join(" " + cxx_float(v) for v in vals[:10])]


# ---- V4 (ref v4_mpi_cuda/src/main_mpi_cuda.cpp:133-136,157-159) -----------------------------------
def v4_shape_values(H: int, W: int, C: int, vals: Sequence[float]):
    return [f"Final Output Shape: {H}x{W}x{C}",
            "Final Output (first 10 values): " + " ".join(cxx_float(v) for v in vals[:10])]


def v4_time(ms: float, label: str = "MPI+CUDA") -> str:
    return f"AlexNet {label} Forward Pass completed in {cxx_float(ms)} ms"


# ---- compatible additions (printed AFTER the canonical lines; never "<num> ms") -------------------
def extras(shape: Optional[Sequence[int]] = None, fnv: Optional[int] = None,
           wall_ms: Optional[float] = None, **kv) -> Iterable[str]:
    if shape is not None:
        yield "Final Output Shape: " + "x".join(str(int(s)) for s in shape)
    if fnv is not None:
        yield f"Output FNV1a64: 0x{fnv:016x}"
    for k, v in kv.items():
        if v is not None:
            yield f"{k}={cxx_float(v) if isinstance(v, float) else v}"
    if wall_ms is not None:
        yield f"ExecutionTime_ms: {cxx_float(wall_ms)}"     # populates log_analysis' run_logs table


def rank_err(rank: int, msg: str) -> None:
    """``[Rank r] ...`` diagnostics on stderr (ref v4_mpi_cuda/src/alexnet_mpi_cuda.cu:66,83)."""
    print(f"[Rank {rank}] {msg}", file=sys.stderr, flush=True)
