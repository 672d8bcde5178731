"""Real multi-GPU runs (one process per GPU over NCCL bootstrap + CUDA IPC peer memory).
Needs >= 2 GPUs: spawned through bin/mpirun like the reference's `mpirun -np N ./template`."""
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _run(np_, variant, extra=()):
    cmd = [sys.executable, os.path.join(ROOT, "bin", "mpirun"), "--oversubscribe", "-np", str(np_),
           sys.executable, "-m", "b200cnn.runners.v4", variant, "--init", "rand", "--seed", "3",
           "--batch", "4", "--iters", "3", *extra]
    env = dict(os.environ, PYTHONPATH=ROOT)
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert p.returncode == 0, p.stdout + p.stderr
    fnv = re.search(r"Output FNV1a64: (0x[0-9a-f]+)", p.stdout).group(1)
    shape = re.search(r"^Final Output Shape: (\S+)", p.stdout, re.M).group(1)
    return fnv, shape, p.stdout


@pytest.mark.parametrize("variant", ["v4", "v5"])
def test_np2_equals_np1(variant):
    f1, s1, _ = _run(1, variant)
    f2, s2, out = _run(2, variant)
    assert s1 == s2 == "52x13x256"          # batch 4 x 13 rows
    assert f1 == f2, out


def test_all_gpus_equal_np1():
    n = torch.cuda.device_count()
    f1, _, _ = _run(1, "v4")
    fn, _, out = _run(n, "v4")
    assert f1 == fn, out


@pytest.mark.parametrize("extra", [("--gather", "host"), ("--row-groups", "auto", "--gather", "host"),
                                   ("--row-groups", "auto", "--strategy", "recompute")])
def test_all_gpus_rows_x_batch_and_host_gather_equal_np1(extra):
    """Rows x batch decomposition, the barrier-free forward and the shared-host-tensor gather on real peers."""
    n = torch.cuda.device_count()
    f1, _, _ = _run(1, "v5")
    fn, _, out = _run(n, "v5", extra)
    assert f1 == fn, out


def _run_full(np_, extra=()):
    cmd = [sys.executable, os.path.join(ROOT, "bin", "mpirun"), "--oversubscribe", "-np", str(np_),
           sys.executable, "-m", "b200cnn.runners.full", "--batch", "8", "--seed", "3", *extra]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, PYTHONPATH=ROOT), cwd=ROOT)
    assert p.returncode == 0, p.stdout + p.stderr
    vals = [float(v) for v in re.search(r"^Final Output \(first 10 values\): (.*)$", p.stdout, re.M).group(1).split()]
    am = re.search(r"argmax0=(\d+)", p.stdout).group(1)
    fnv = re.search(r"Output FNV1a64: (0x[0-9a-f]+)", p.stdout).group(1)
    return vals, am, fnv


@pytest.mark.parametrize("extra", [(), ("--row-groups", "auto")])
def test_full_network_all_gpus_equals_np1_and_is_deterministic(extra):
    """Full AlexNet: decomposed blocks 1-2, batch-sharded conv3-5, tensor-parallel FC over real peer memory.  The
    FC7 reduce-scatter adds the ranks' partial products in rank order, so two runs give the same bits; against one
    rank only the fp32 summation order of FC7/FC8 differs."""
    n = torch.cuda.device_count()
    v1, a1, _ = _run_full(1)
    vn, an, fn = _run_full(n, extra)
    _, _, fn2 = _run_full(n, extra)
    assert fn == fn2, "two runs of the same np differ"
    assert a1 == an
    for a, b in zip(v1, vn):
        assert abs(a - b) <= 1e-4 * max(abs(a), 1e-6) + 1e-8
