"""Native V2 `template` binaries (csrc/apps/v2_main.cpp: native planner + slab layers + MPI subset) built by the
version Makefiles: np N prints what np 1 and the serial V1 binary print (exact decomposition), for both variants,
both strategies, a batch and a non-square input; bad arguments fail-stop with a [Rank r] line."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FP = os.path.join(ROOT, "final_project")
MPIRUN = [sys.executable, os.path.join(ROOT, "bin", "mpirun"), "--oversubscribe"]


def _make(sub):
    d = os.path.join(FP, sub)
    p = subprocess.run(["make"], cwd=d, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout + p.stderr
    exe = os.path.join(d, "template")
    with open(exe, "rb") as f:
        assert f.read(4) == b"\x7fELF", "V2 template must be a compiled binary"
    return exe


def _run(exe, np_, *args):
    p = subprocess.run(MPIRUN + ["-np", str(np_), exe, *args], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    fnv = re.search(r"Output FNV1a64: (0x[0-9a-f]+)", p.stdout).group(1)
    shape = re.search(r"^shape: (\S+)", p.stdout, re.M).group(1)
    assert re.search(r"^Sample values: \S+ \S+ \S+ \S+ \S+$", p.stdout, re.M)
    assert re.search(r"^Execution Time: [0-9.e+-]+ ms$", p.stdout, re.M)
    return shape, fnv


@pytest.fixture(scope="module")
def exes():
    return {"v1": _make("v1_serial"), "v21": _make("v2_mpi_only/2.1_broadcast_all"),
            "v22": _make("v2_mpi_only/2.2_scatter_halo")}


def test_native_v22_equals_serial_v1_at_every_np(exes):
    v1 = subprocess.run([exes["v1"], "--seed", "11"], capture_output=True, text=True, timeout=300).stdout
    want = re.search(r"Output FNV1a64: (0x[0-9a-f]+)", v1).group(1)
    for np_ in (1, 2, 3, 5, 8, 16):          # 16 ranks for 13 final rows: three ranks own nothing
        assert _run(exes["v22"], np_, "--init", "rand", "--seed", "11") == ("13x13x256", want), np_
    assert _run(exes["v21"], 3, "--init", "rand", "--seed", "11") == ("13x13x256", want)


def test_native_v22_strategies_batch_and_shapes(exes):
    base = ("--height", "99", "--width", "131", "--batch", "3", "--init", "rand", "--seed", "5", "--lrn", "alpha_raw")
    ref = _run(exes["v22"], 1, *base)
    assert ref[0] == "5x7x256"
    for np_, strat in ((2, "halo"), (4, "halo"), (7, "halo"), (3, "recompute"), (6, "recompute")):
        assert _run(exes["v22"], np_, *base, "--strategy", strat) == ref, (np_, strat)
    # the Python runner over gloo prints the same checksum: one planner, two hosts for it
    env = dict(os.environ, PYTHONPATH=ROOT)
    p = subprocess.run(MPIRUN + ["-np", "2", sys.executable, "-m", "b200cnn.runners.v2_2", *base], env=env,
                       capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr
    assert re.search(r"Output FNV1a64: (0x[0-9a-f]+)", p.stdout).group(1) == ref[1]


def test_native_v2_constant_init_matches_the_closed_form_fixture(exes):
    # constant init, alpha/N LRN -> first values 44.4152 42.4612 40.6967 ...
    p = subprocess.run(MPIRUN + ["-np", "4", exes["v22"]], capture_output=True, text=True, timeout=300)
    assert "Sample values: 44.4152 42.4612 40.6967 40.6967 40.6967" in p.stdout


def test_native_v2_fail_stop(exes):
    p = subprocess.run(MPIRUN + ["-np", "2", exes["v22"], "--no-such-flag"], capture_output=True, text=True, timeout=60)
    assert p.returncode != 0 and "[Rank" in p.stderr and "unknown argument --no-such-flag" in p.stderr
