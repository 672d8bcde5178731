"""Homework toolchain: scaffold / test / package / run (ref scripts/scaffold_hw.sh, test_hw.sh,
package_hw.sh, run_hw.sh).

* ``scaffold N``  creates ``homeworks/hwN/`` (Makefile, src/template.*, summary.md) from templates/;
                  HW1-3 get the distributed skeleton, HW4+ additionally the CUDA + CMake skeleton
                  (ref scaffold_hw.sh:475-520).
* ``test N``      builds, then runs np in PROCESS_COUNTS x n in PROBLEM_SIZES, skipping n % np != 0,
                  each under TIMEOUT_DURATION; a run passes when it exits 0 and prints ``Test: PASSED``.
                  Exit code 0 = all passed, 1 = a failure, 2 = only timeouts (inconclusive)
                  (ref test_hw.sh:5-10,113-180).
* ``package N``   ``hwN-<last>-<first>.tgz`` with ``template.*`` and the Makefile (ref package_hw.sh:18-19,82).
* ``run N``       test, then package if the tests passed (ref run_hw.sh).
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
import tarfile

from .. import REPO_ROOT

PROBLEM_SIZES = [128, 256, 512, 1024, 2048]
PROCESS_COUNTS = [1, 2, 3, 4, 5, 6, 7, 8]
TIMEOUT_DURATION = 30


def hw_dir(n: int) -> str:
    return os.path.join(REPO_ROOT, "homeworks", f"hw{n}")


def _render(name: str, hw: str) -> str:
    with open(os.path.join(REPO_ROOT, "templates", name)) as f:
        return f.read().replace("{{HW}}", hw)


def scaffold(n: int, force: bool = False, native: bool = False) -> int:
    """``native``: the compiled C skeleton (templates/template.c.template + Makefile.native.template, MPI subset of
    baseline/mpishim) instead of the Python one — the counterpart of ref templates/template.c.template."""
    d = hw_dir(n)
    hw = f"hw{n}"
    if os.path.exists(d) and not force:
        print(f"{d} exists (use --force to overwrite)")
        return 1
    os.makedirs(os.path.join(d, "src"), exist_ok=True)
    files = {"Makefile": _render("Makefile.native.template" if native else "Makefile.template", hw),
             ("src/template.c" if native else "src/template.py"):
                 _render("template.c.template" if native else "template.py.template", hw),
             "summary.md": f"# {hw.upper()} summary\n\n- approach:\n- results:\n- issues:\n"}
    if n >= 4:
        files["src/template.cu"] = _render("template.cu.template", hw)
        files["CMakeLists.txt"] = _render("CMakeLists.txt.template", hw)
    for rel, text in files.items():
        with open(os.path.join(d, rel), "w") as f:
            f.write(text)
    print(f"scaffolded {d}: {', '.join(sorted(files))}")
    return 0


def test(n: int, sizes=None, counts=None, timeout=None) -> int:
    d = hw_dir(n)
    sizes, counts = sizes or PROBLEM_SIZES, counts or PROCESS_COUNTS
    timeout = timeout or TIMEOUT_DURATION
    if subprocess.run(["make", "clean"], cwd=d, capture_output=True).returncode or \
            subprocess.run(["make"], cwd=d, capture_output=True).returncode:
        print("build FAILED")
        return 1
    mpirun = os.path.join(REPO_ROOT, "bin", "mpirun")
    npass = nfail = ntime = nskip = 0
    for np_ in counts:
        for sz in sizes:
            if sz % np_ != 0:
                nskip += 1
                continue
            cmd = [mpirun, "--oversubscribe", "-np", str(np_), "./template", str(sz)]
            try:
                p = subprocess.run(cmd, cwd=d, capture_output=True, text=True, timeout=timeout)
                ok = p.returncode == 0 and "Test: PASSED" in p.stdout
                status = "PASS" if ok else f"FAIL (rc={p.returncode})"
                npass += ok
                nfail += not ok
            except subprocess.TimeoutExpired:
                status = f"TIMEOUT (> {timeout}s, exit 124)"
                ntime += 1
            print(f"np={np_:<2} n={sz:<5} {status}", flush=True)
    print(f"passed={npass} failed={nfail} timeouts={ntime} skipped={nskip}")
    return 1 if nfail else (2 if ntime else 0)


def package(n: int, last: str, first: str) -> int:
    d = hw_dir(n)
    out = os.path.join(REPO_ROOT, f"hw{n}-{last.lower()}-{first.lower()}.tgz")
    members = [os.path.join(d, "Makefile")]
    src = os.path.join(d, "src")
    members += [os.path.join(src, f) for f in sorted(os.listdir(src)) if f.startswith("template.")]
    with tarfile.open(out, "w:gz") as t:
        for m in members:
            if os.path.exists(m):
                t.add(m, arcname=os.path.join(f"hw{n}", os.path.relpath(m, d)))
    print(f"packaged {len(members)} files -> {out}")
    return 0


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="hw")
    sub = ap.add_subparsers(dest="cmd", required=True)
    for c in ("scaffold", "test", "package", "run"):
        a = sub.add_parser(c)
        a.add_argument("n", type=int)
        if c == "scaffold":
            a.add_argument("--force", action="store_true")
            a.add_argument("--native", action="store_true",
                           help="compiled C skeleton (MPI subset of baseline/mpishim) instead of the Python one")
        if c in ("test", "run"):
            a.add_argument("--sizes", default="")
            a.add_argument("--counts", default="")
            a.add_argument("--timeout", type=int, default=TIMEOUT_DURATION)
        if c in ("package", "run"):
            a.add_argument("--last", default=os.environ.get("HW_LASTNAME", "lastname"))
            a.add_argument("--first", default=os.environ.get("HW_FIRSTNAME", "firstname"))
    ns = ap.parse_args(argv)
    ints = lambda s: [int(x) for x in s.split(",") if x]
    if ns.cmd == "scaffold":
        return scaffold(ns.n, ns.force, ns.native)
    if ns.cmd == "test":
        return test(ns.n, ints(ns.sizes), ints(ns.counts), ns.timeout)
    if ns.cmd == "package":
        return package(ns.n, ns.last, ns.first)
    rc = test(ns.n, ints(ns.sizes), ints(ns.counts), ns.timeout)
    if rc != 0:
        print("tests did not pass; not packaging")
        return rc
    return package(ns.n, ns.last, ns.first)


if __name__ == "__main__":
    sys.exit(main())
