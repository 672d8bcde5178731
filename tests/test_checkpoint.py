"""Weight files (utils/checkpoint.py): round trip, memory-mapped shard access, corruption and wrong-shape refusal,
resume-or-initialise, and the runners' --save-weights / --load-weights flags (CPU)."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem, Blocks12Weights
from b200cnn.utils import checkpoint as ck

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _cfg(**kw):
    return default_config(H=63, W=63, batch=2, init=INIT_RAND, seed=3, **kw)


def test_blocks12_round_trip_is_bit_exact(tmp_path):
    cfg = _cfg()
    _, w = make_problem(cfg)
    path = str(tmp_path / "w.b2w")
    header = ck.save_weights(path, w, meta={"seed": 3})
    assert header["kind"] == "blocks12" and [e["name"] for e in header["tensors"]] == ["w1", "b1", "w2", "b2"]
    assert all(e["offset"] % 256 == 0 for e in header["tensors"])
    r = ck.load_weights(path)
    assert isinstance(r, Blocks12Weights)
    for k in ("w1", "b1", "w2", "b2"):
        assert torch.equal(getattr(r, k), getattr(w, k))
    assert ck.read_header(path)["meta"] == {"seed": 3}
    assert not [f for f in os.listdir(tmp_path) if ".tmp." in f]          # written atomically


def test_full_weights_mmap_shard(tmp_path):
    from b200cnn.models.alexnet_full import make_full_weights, FullWeights
    cfg = _cfg()
    _, w12 = make_problem(cfg)
    w = make_full_weights(w12, seed=5, feat=256)
    path = str(tmp_path / "full.b2w")
    ck.save_weights(path, w)
    m = ck.load_weights(path, mmap=True)
    assert isinstance(m, FullWeights)
    # a tensor-parallel rank's FC6 shard: rows [1024, 2048) — a view over the mapping, no copy of the file
    shard = m.w6[1024:2048]
    assert torch.equal(shard, w.w6[1024:2048])
    assert torch.equal(m.w8, w.w8) and torch.equal(m.blocks12.w2, w12.w2)


def test_corruption_truncation_and_wrong_shape_are_refused(tmp_path):
    cfg = _cfg()
    _, w = make_problem(cfg)
    path = str(tmp_path / "w.b2w")
    ck.save_weights(path, w)
    h = ck.read_header(path)
    raw = bytearray(open(path, "rb").read())
    pos = h["_data_start"] + h["tensors"][2]["offset"] + 40
    raw[pos] ^= 0x01
    bad = str(tmp_path / "bad.b2w")
    open(bad, "wb").write(raw)
    with pytest.raises(ValueError, match="checksum of w2"):
        ck.load_weights(bad)
    raw[pos] ^= 0x01
    open(bad, "wb").write(bytes(raw[:len(raw) - 64]))
    with pytest.raises(ValueError, match="truncated"):
        ck.load_weights(bad)
    open(bad, "wb").write(b"not a weight file at all")
    with pytest.raises(ValueError, match="not a b200cnn weight file"):
        ck.load_weights(bad)
    other = default_config(H=63, W=63, C=4, batch=2, init=INIT_RAND, seed=3)      # a 4-channel input
    with pytest.raises(ValueError, match="w1 has shape"):
        ck.check_against(other, ck.load_weights(path))


def test_resume_or_init(tmp_path):
    cfg = _cfg()
    path = str(tmp_path / "resume.b2w")
    calls = []

    def make():
        calls.append(1)
        return make_problem(cfg)[1]
    w0, how0 = ck.resume_or_init(cfg, path, make, save_if_missing=True)
    w1, how1 = ck.resume_or_init(cfg, path, make, save_if_missing=True)
    assert (how0, how1) == ("initialised+saved", "loaded") and len(calls) == 1
    assert torch.equal(w0.w2, w1.w2)
    w2, how2 = ck.resume_or_init(cfg, None, make)
    assert how2 == "initialised" and len(calls) == 2


def test_v1_runner_saves_and_loads_weights(tmp_path):
    """`--save-weights` writes what the run used; `--load-weights` of that file reproduces the output checksum even
    under another seed (the input stream is seeded separately from the loaded parameters only through the file)."""
    env = dict(os.environ, PYTHONPATH=ROOT)
    path = str(tmp_path / "v1.b2w")
    base = [sys.executable, "-m", "b200cnn.runners.v1", "--height", "63", "--width", "63", "--init", "rand", "--seed", "7"]
    a = subprocess.run(base + ["--save-weights", path], env=env, capture_output=True, text=True, timeout=300)
    assert a.returncode == 0, a.stderr
    assert os.path.exists(path)
    b = subprocess.run(base + ["--load-weights", path], env=env, capture_output=True, text=True, timeout=300)
    assert b.returncode == 0, b.stderr
    fa = [l for l in a.stdout.splitlines() if "FNV1a64" in l]
    fb = [l for l in b.stdout.splitlines() if "FNV1a64" in l]
    assert fa and fa == fb
    c = subprocess.run(base + ["--load-weights", str(tmp_path / "missing.b2w")], env=env, capture_output=True,
                       text=True, timeout=300)
    assert c.returncode != 0 and "missing.b2w" in (c.stderr + c.stdout)


def test_native_reader_and_native_binaries_load_the_same_file(tmp_path):
    """b200_weights_read (csrc/cpu/weights_cpu.cpp) returns what Python wrote, refuses a flipped bit, and the native
    V1 / V2.2 binaries started with --load-weights print the checksum of the Python runner given the same file."""
    from b200cnn.ops.cpu import native_weights_read
    cfg = default_config(batch=1, init=INIT_RAND, seed=21)
    _, w = make_problem(cfg)
    path = str(tmp_path / "w.b2w")
    ck.save_weights(path, w)
    for k in ("w1", "b1", "w2", "b2"):
        assert torch.equal(native_weights_read(path, k), getattr(w, k))
    with pytest.raises(ValueError):
        native_weights_read(path, "w9")
    h = ck.read_header(path)
    raw = bytearray(open(path, "rb").read())
    raw[h["_data_start"] + h["tensors"][0]["offset"] + 8] ^= 0x10
    bad = str(tmp_path / "bad.b2w")
    open(bad, "wb").write(raw)
    with pytest.raises(ValueError):
        native_weights_read(bad, "w1")
    # native binaries: V1 (seed 3 input, parameters from the file) == native V2.2 np 3 == Python V1 runner
    fp = os.path.join(ROOT, "final_project")
    for sub in ("v1_serial", "v2_mpi_only/2.2_scatter_halo"):
        assert subprocess.run(["make"], cwd=os.path.join(fp, sub), capture_output=True, timeout=600).returncode == 0

    def fnv(out):
        return [l for l in out.splitlines() if "FNV1a64" in l]
    a = subprocess.run([os.path.join(fp, "v1_serial", "template"), "--seed", "3", "--load-weights", path],
                       capture_output=True, text=True, timeout=300)
    b = subprocess.run([sys.executable, os.path.join(ROOT, "bin", "mpirun"), "--oversubscribe", "-np", "3",
                        os.path.join(fp, "v2_mpi_only/2.2_scatter_halo", "template"), "--init", "rand", "--seed", "3",
                        "--load-weights", path], capture_output=True, text=True, timeout=300)
    c = subprocess.run([sys.executable, "-m", "b200cnn.runners.v1", "--init", "rand", "--seed", "3", "--load-weights", path],
                       env=dict(os.environ, PYTHONPATH=ROOT), capture_output=True, text=True, timeout=300)
    assert a.returncode == 0 and b.returncode == 0 and c.returncode == 0, (a.stderr, b.stderr, c.stderr)
    assert fnv(a.stdout) and fnv(a.stdout) == fnv(b.stdout) == fnv(c.stdout)
    d = subprocess.run([os.path.join(fp, "v1_serial", "template"), "--load-weights", bad], capture_output=True, text=True)
    assert d.returncode != 0 and "--load-weights" in d.stderr
