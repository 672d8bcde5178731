"""Build-run-parse harness: the equivalent of the reference's bash test scripts.

What it reproduces (ref scripts/common_test_utils.sh:13-346, scripts/0_run_final_project.sh:44-73,
scripts/1_final_unique_machine.sh:43-72, final_project/v4_mpi_cuda/test_v4.sh:82-123):

* GPU arch detection -> ``GPU_ARCH_FLAGS`` handed to ``make`` for CUDA variants;
* per case: ``make clean && make`` in the variant directory, then ``./template`` or
  ``mpirun --oversubscribe -np N ./template`` (variants whose name contains V2/V4/V5);
* failure classification by exit code + log grep (0 ok / 2 environment / 3 MPI config /
  4 segfault / 124 timeout / 1 other);
* regex parsing of the run log: time = FIRST ``<num> ms``; shape = ``Final Output Shape:`` else
  first ``Dimensions:`` else ``shape:``; first values = ``Final Output (first 10 values):`` else
  ``Sample values:``;
* one CSV row per case with the reference's 20-column header, an ASCII summary table, and a
  session directory ``final_project/logs/<prefix>_session_<ts>_<host>/`` holding
  ``make_<tag>_npN.log`` / ``run_<tag>_npN.log`` / ``summary_report_<ts>.csv``.
"""
from __future__ import annotations

import argparse
import csv
import datetime
import os
import re
import shutil
import socket
import subprocess
import sys
from dataclasses import dataclass
from typing import List, Optional, Tuple

from .. import REPO_ROOT

CSV_HEADER = ("SessionID,MachineID,GitCommit,EntryTimestamp,ProjectVariant,NumProcesses,MakeLogFile,"
              "BuildSucceeded,BuildMessage,RunLogFile,RunCommandSucceeded,RunEnvironmentWarning,"
              "RunMessage,ParseSucceeded,ParseMessage,OverallStatusSymbol,OverallStatusMessage,"
              "ExecutionTime_ms,OutputShape,OutputFirst5Values").split(",")

FINAL = os.path.join(REPO_ROOT, "final_project")


@dataclass
class Case:
    name: str            # display name, e.g. "V2 2.2-scatter-halo"
    subdir: str          # directory under final_project/
    tag: str             # log file tag
    np: int
    cuda: bool = False
    mpi: bool = False


def default_cases(np_values=(1, 2, 4), with_v5: bool = True) -> List[Case]:
    cs = [Case("V1 Serial", "v1_serial", "v1_serial", 1)]
    for sub, nm in (("2.1_broadcast_all", "V2 2.1-broadcast-all"), ("2.2_scatter_halo", "V2 2.2-scatter-halo")):
        for n in np_values:
            cs.append(Case(nm, f"v2_mpi_only/{sub}", "v2_" + sub.replace(".", "_"), n, mpi=True))
    cs.append(Case("V3 CUDA Only", "v3_cuda_only", "v3_cuda_only", 1, cuda=True))
    for n in np_values:
        cs.append(Case("V4 MPI+CUDA", "v4_mpi_cuda", "v4_mpi_cuda", n, cuda=True, mpi=True))
    if with_v5:
        for n in np_values:
            cs.append(Case("V5 MPI+CUDA-Aware", "v5_cuda_aware_mpi", "v5_cuda_aware_mpi", n, cuda=True, mpi=True))
    return cs


# ---- GPU arch detection (ref common_test_utils.sh:13-68) -------------------------------------
def detect_gpu_arch_flags() -> Tuple[str, str]:
    """Returns (flags, how).  nvidia-smi compute_cap -> -gencode; without a GPU the build still
    cross-compiles for sm_100a (the only arch the kernels support)."""
    env = os.environ.get("DETECTED_GPU_ARCH_FLAGS")
    if env:
        return env, "DETECTED_GPU_ARCH_FLAGS from the environment"
    smi = shutil.which("nvidia-smi")
    if smi:
        try:
            out = subprocess.run([smi, "--query-gpu=compute_cap", "--format=csv,noheader"],
                                 capture_output=True, text=True, timeout=20).stdout.strip().splitlines()
            if out:
                cc = out[0].strip().replace(".", "")
                suffix = "a" if cc in ("90", "100", "101", "103", "120") else ""
                return (f"-gencode arch=compute_{cc}{suffix},code=sm_{cc}{suffix}",
                        f"nvidia-smi compute_cap {out[0].strip()}")
        except Exception:
            pass
    return "-gencode arch=compute_100a,code=sm_100a", "no GPU visible: default sm_100a (cross-compile)"


# ---- log parsing (ref common_test_utils.sh:296-317) --------------------------------------------
RE_TIME = re.compile(r"([0-9]+(?:\.[0-9]+)?) ms")
RE_SHAPE_FINAL = re.compile(r"^Final Output Shape:\s*(\d+x\d+x\d+)", re.M)
RE_SHAPE_DIMS = re.compile(r"Dimensions: H=(\d+), W=(\d+), C=(\d+)")
RE_SHAPE_V2 = re.compile(r"^shape:\s*(\d+x\d+x\d+)", re.M)
RE_VALS10 = re.compile(r"^Final Output \(first 10 values\):\s*(.*)$", re.M)
RE_VALS5 = re.compile(r"^Sample values:\s*(.*)$", re.M)


def parse_run_log(text: str) -> Tuple[Optional[str], Optional[str], Optional[str]]:
    m = RE_TIME.search(text)
    t = m.group(1) if m else None
    shape = None
    m = RE_SHAPE_FINAL.search(text)
    if m:
        shape = m.group(1)
    else:
        m = RE_SHAPE_DIMS.search(text)
        if m:
            shape = "x".join(m.groups())
        else:
            m = RE_SHAPE_V2.search(text)
            if m:
                shape = m.group(1)
    vals = None
    m = RE_VALS10.search(text) or RE_VALS5.search(text)
    if m:
        toks = m.group(1).replace("...", " ").split()
        vals = " ".join(toks[:5])
    return t, shape, vals


def classify_failure(rc: int, log: str) -> Tuple[int, str]:
    """ref common_test_utils.sh:84-117."""
    low = log.lower()
    if rc == 124:
        return 124, "timeout"
    if "no cuda-capable device" in low or "driver version is insufficient" in low or \
            "cuda_error_no_device" in low:
        return 2, "environment: no usable CUDA device"
    if "not enough slots" in low or "slots available" in low or "there are not enough slots" in low:
        return 3, "MPI configuration: slots unavailable"
    if rc in (139, -11) or "segmentation fault" in low:
        return 4, "segmentation fault"
    return 1, f"exit code {rc}"


def git_commit() -> str:
    try:
        return subprocess.run(["git", "-C", REPO_ROOT, "rev-parse", "--short", "HEAD"],
                              capture_output=True, text=True, timeout=10).stdout.strip() or "unknown"
    except Exception:
        return "unknown"


def _run(cmd, cwd, logfile, timeout, env=None) -> int:
    with open(logfile, "w") as f:
        f.write(f"$ {' '.join(cmd)}\n")
        f.flush()
        try:
            return subprocess.run(cmd, cwd=cwd, stdout=f, stderr=subprocess.STDOUT, timeout=timeout,
                                  env=env).returncode
        except subprocess.TimeoutExpired:
            f.write("\n[harness] timeout\n")
            return 124


def execute_case(c: Case, session_dir: str, session_id: str, arch_flags: str, timeout: int,
                 extra_args: List[str]) -> dict:
    vdir = os.path.join(FINAL, c.subdir)
    mk = os.path.join(session_dir, f"make_{c.tag}_np{c.np}.log")
    rl = os.path.join(session_dir, f"run_{c.tag}_np{c.np}.log")
    row = {h: "" for h in CSV_HEADER}
    row.update(SessionID=session_id, MachineID=socket.gethostname(), GitCommit=git_commit(),
               EntryTimestamp=datetime.datetime.now().isoformat(timespec="seconds"),
               ProjectVariant=c.name, NumProcesses=c.np, MakeLogFile=os.path.basename(mk),
               RunLogFile=os.path.basename(rl))
    env = dict(os.environ)
    env["PATH"] = os.path.join(REPO_ROOT, "bin") + os.pathsep + env.get("PATH", "")
    # -- build: make clean && make [GPU_ARCH_FLAGS=...] --------------------------------------------
    make = ["make"] + ([f"GPU_ARCH_FLAGS={arch_flags}"] if c.cuda else [])
    rc = _run(["make", "clean"], vdir, mk, 120, env)
    with open(mk, "a") as f:
        rc = subprocess.run(make, cwd=vdir, stdout=f, stderr=subprocess.STDOUT, env=env).returncode
    ok_build = rc == 0 and os.access(os.path.join(vdir, "template"), os.X_OK)
    row["BuildSucceeded"] = "true" if ok_build else "false"
    row["BuildMessage"] = "Build OK" if ok_build else f"make failed (rc={rc})"
    if not ok_build:
        row.update(RunCommandSucceeded="false", ParseSucceeded="false", OverallStatusSymbol="✘",
                   OverallStatusMessage="Build failed")
        return row
    # -- run ----------------------------------------------------------------------------------------
    if c.mpi:
        cmd = ["mpirun", "--oversubscribe", "-np", str(c.np), "./template"] + extra_args
    else:
        cmd = ["./template"] + extra_args
    rc = _run(cmd, vdir, rl, timeout, env)
    text = open(rl).read()
    if rc == 0:
        row.update(RunCommandSucceeded="true", RunEnvironmentWarning="false", RunMessage="Run OK")
    else:
        code, msg = classify_failure(rc, text)
        row.update(RunCommandSucceeded="false", RunEnvironmentWarning="true" if code in (2, 3) else "false",
                   RunMessage=msg)
    # -- parse ----------------------------------------------------------------------------------------
    t, shape, vals = parse_run_log(text.split("\n", 1)[1] if "\n" in text else text)
    parsed = rc == 0 and t is not None and shape is not None and vals is not None
    row["ParseSucceeded"] = "true" if parsed else "false"
    row["ParseMessage"] = "Parsed OK" if parsed else "missing: " + ",".join(
        n for n, v in (("time", t), ("shape", shape), ("values", vals)) if v is None)
    row.update(ExecutionTime_ms=t or "", OutputShape=shape or "", OutputFirst5Values=vals or "")
    if rc == 0 and parsed:
        row.update(OverallStatusSymbol="✔", OverallStatusMessage="OK")
    elif row["RunEnvironmentWarning"] == "true":
        row.update(OverallStatusSymbol="⚠", OverallStatusMessage=row["RunMessage"])
    else:
        row.update(OverallStatusSymbol="✘", OverallStatusMessage=row["RunMessage"] if rc else row["ParseMessage"])
    return row


def summary_table(rows: List[dict]) -> str:
    cols = [("ProjectVariant", 22), ("NumProcesses", 3), ("OverallStatusSymbol", 2),
            ("ExecutionTime_ms", 12), ("OutputShape", 12), ("OutputFirst5Values", 44),
            ("OverallStatusMessage", 30)]
    head = ("Variant", "NP", "St", "Time (ms)", "Shape", "First 5 values", "Message")
    line = "+" + "+".join("-" * (w + 2) for _, w in cols) + "+"
    out = [line, "| " + " | ".join(h.ljust(w) for h, (_, w) in zip(head, cols)) + " |", line]
    for r in rows:
        out.append("| " + " | ".join(str(r[k])[:w].ljust(w) for k, w in cols) + " |")
    out.append(line)
    return "\n".join(out)


def run_suite(prefix: str, cases: List[Case], timeout: int = 600, extra_args: Optional[List[str]] = None,
              logs_root: Optional[str] = None) -> Tuple[str, List[dict]]:
    ts = datetime.datetime.now().strftime("%Y%m%d_%H%M%S")
    host = socket.gethostname()
    session_id = f"{prefix}_session_{ts}_{host}"
    session_dir = os.path.join(logs_root or os.path.join(FINAL, "logs"), session_id)
    os.makedirs(session_dir, exist_ok=True)
    flags, how = detect_gpu_arch_flags()
    print(f"[harness] session {session_id}\n[harness] GPU arch flags: {flags} ({how})", flush=True)
    rows = []
    csv_path = os.path.join(session_dir, f"summary_report_{ts}.csv")
    with open(csv_path, "w", newline="") as f:
        wtr = csv.DictWriter(f, fieldnames=CSV_HEADER)
        wtr.writeheader()
        for c in cases:
            print(f"[harness] {c.name} np={c.np} ...", flush=True)
            row = execute_case(c, session_dir, session_id, flags, timeout, extra_args or [])
            rows.append(row)
            wtr.writerow(row)
            f.flush()
    print(summary_table(rows))
    print(f"[harness] CSV: {csv_path}")
    return csv_path, rows


# ---- single-case entry points for scripts/common_test_utils.sh --------------------------------
def run_one_case(name: str, np_: int, subdir: str, tag: str, session_dir: str, csv_path: str,
                 timeout: int = 600, extra_args: Optional[List[str]] = None) -> dict:
    """One build-run-parse case appended to ``csv_path`` (created with the header when missing): what the
    reference's ``setup_test_case`` + ``execute_single_test_case`` pair does (ref
    scripts/common_test_utils.sh:187-346).  A missing variant directory is logged as a failed row."""
    up = name.upper()
    c = Case(name, subdir, tag, np_, cuda=any(k in up for k in ("V3", "V4", "V5", "CUDA")),
             mpi=any(k in up for k in ("V2", "V4", "V5", "MPI")))
    os.makedirs(session_dir, exist_ok=True)
    session_id = os.path.basename(os.path.normpath(session_dir))
    if not os.path.isdir(os.path.join(FINAL, subdir)):
        row = {h: "" for h in CSV_HEADER}
        row.update(SessionID=session_id, MachineID=socket.gethostname(), GitCommit=git_commit(),
                   EntryTimestamp=datetime.datetime.now().isoformat(timespec="seconds"), ProjectVariant=name,
                   NumProcesses=np_, MakeLogFile="N/A", BuildSucceeded="false", BuildMessage="Dir not found",
                   RunLogFile="N/A", RunCommandSucceeded="false", RunEnvironmentWarning="false",
                   RunMessage="Skipped", ParseSucceeded="false", ParseMessage="Skipped",
                   OverallStatusSymbol="✘", OverallStatusMessage="✘ Dir Missing")
    else:
        flags = os.environ.get("DETECTED_GPU_ARCH_FLAGS") or detect_gpu_arch_flags()[0]
        row = execute_case(c, session_dir, session_id, flags, timeout, extra_args or [])
    new = not os.path.exists(csv_path)
    with open(csv_path, "a", newline="") as f:
        wtr = csv.DictWriter(f, fieldnames=CSV_HEADER)
        if new:
            wtr.writeheader()
        wtr.writerow(row)
    return row


def _sub_main(argv: List[str]) -> Optional[int]:
    """``arch`` | ``case ...`` | ``table <csv>``: the pieces a bash caller composes itself."""
    if not argv or argv[0] not in ("arch", "case", "table"):
        return None
    if argv[0] == "arch":
        flags, how = detect_gpu_arch_flags()
        print(flags)
        print(f"[ARCH_DETECT] {how}", file=sys.stderr)
        return 0
    if argv[0] == "table":
        with open(argv[1], newline="") as f:
            print(summary_table(list(csv.DictReader(f))))
        return 0
    ap = argparse.ArgumentParser(prog="suite case")
    ap.add_argument("--name", required=True); ap.add_argument("--np", type=int, default=1)
    ap.add_argument("--subdir", required=True); ap.add_argument("--tag", required=True)
    ap.add_argument("--session-dir", required=True); ap.add_argument("--csv", required=True)
    ap.add_argument("--timeout", type=int, default=600)
    ap.add_argument("rest", nargs="*")
    ns = ap.parse_args(argv[1:])
    row = run_one_case(ns.name, ns.np, ns.subdir, ns.tag, ns.session_dir, ns.csv, ns.timeout, ns.rest)
    print(f"  [{row['OverallStatusSymbol']}] {ns.name} np={ns.np}: {row['OverallStatusMessage']}"
          f"  time={row['ExecutionTime_ms'] or '-'} ms  shape={row['OutputShape'] or '-'}")
    return 0 if row["OverallStatusSymbol"] != "✘" else 1


def main(argv=None) -> int:
    sub = _sub_main(list(sys.argv[1:] if argv is None else argv))
    if sub is not None:
        return sub
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--session-prefix", default="script1")
    ap.add_argument("--np", default="1,2,4", help="process counts for the MPI variants")
    ap.add_argument("--only", default="", help="comma list of substrings; run matching variants only")
    ap.add_argument("--cpu-only", action="store_true", help="skip CUDA variants (no GPU here)")
    ap.add_argument("--timeout", type=int, default=600)
    ap.add_argument("--logs-root", default=None)
    ap.add_argument("rest", nargs="*", help="extra args passed to every ./template")
    ns = ap.parse_args(argv)
    cases = default_cases(tuple(int(x) for x in ns.np.split(",")))
    if ns.cpu_only:
        cases = [c for c in cases if not c.cuda]
    if ns.only:
        keys = [k.strip().lower() for k in ns.only.split(",")]
        cases = [c for c in cases if any(k in c.name.lower() for k in keys)]
    _, rows = run_suite(ns.session_prefix, cases, ns.timeout, ns.rest, ns.logs_root)
    bad = [r for r in rows if r["OverallStatusSymbol"] == "✘"]
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
