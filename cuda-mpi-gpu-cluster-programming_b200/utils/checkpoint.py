"""Weight files: save / load / resume for blocks 1-2 and the full-network extension.

The reference is stateless (weights are synthesised at start, SURVEY §5.4: "nothing to build beyond optional
weight save/load for the full-network extension").  A framework that serves a trained network needs the other
half, so this is a small, self-describing, memory-mappable container:

    bytes 0..7     magic  b"B200CNNW"
    bytes 8..11    uint32 little-endian: format version (1)
    bytes 12..15   uint32 little-endian: length L of the JSON header
    bytes 16..16+L JSON: {"kind": "blocks12"|"full", "meta": {...}, "tensors": [{"name", "shape", "offset",
                   "nbytes", "fnv1a64"}]}  — offsets are relative to the data section
    data section   starts at the next multiple of 4096; every tensor is raw little-endian float32 in the
                   reference's layouts ([K][C][F][F] convolutions, [out][in] linear layers), 256 B aligned

Every tensor carries its FNV-1a64 (the checksum the runners print for outputs), verified on load.  A rank of a
tensor-parallel job can map the file and touch only its rows (`load_weights(path, mmap=True)` returns views over
a private copy-on-write `numpy.memmap`; pages that are never read are never brought in), and `resume_or_init()` is the one
call a runner makes: load the file if it exists, else synthesise from the seed and (optionally) write it.
"""
from __future__ import annotations

import json
import os
import struct
from typing import Dict, Optional, Union

import numpy as np
import torch

from ..models.init import Blocks12Weights

MAGIC = b"B200CNNW"
VERSION = 1
_DATA_ALIGN = 4096
_TENSOR_ALIGN = 256

_B12 = ("w1", "b1", "w2", "b2")
_FULL = ("w3", "b3", "w4", "b4", "w5", "b5", "w6", "b6", "w7", "b7", "w8", "b8")


def fnv1a64(a: np.ndarray) -> int:
    """FNV-1a64 over the bytes of a contiguous array: the native helper the runners use for their output line."""
    from ..ops.cpu import fnv1a64 as _native
    return _native(torch.from_numpy(np.ascontiguousarray(a)))


def _tensors_of(weights) -> Dict[str, torch.Tensor]:
    if isinstance(weights, Blocks12Weights):
        return {k: getattr(weights, k) for k in _B12}
    out = {k: getattr(weights.blocks12, k) for k in _B12}          # FullWeights
    out.update({k: getattr(weights, k) for k in _FULL})
    return out


def save_weights(path: str, weights, meta: Optional[dict] = None, checksum: bool = True) -> dict:
    """Write ``weights`` (Blocks12Weights or FullWeights) to ``path`` atomically; returns the header."""
    tensors = _tensors_of(weights)
    kind = "blocks12" if isinstance(weights, Blocks12Weights) else "full"
    entries, off = [], 0
    arrays = []
    for name, t in tensors.items():
        a = np.ascontiguousarray(t.detach().to("cpu", torch.float32).numpy()).astype("<f4", copy=False)
        off = (off + _TENSOR_ALIGN - 1) // _TENSOR_ALIGN * _TENSOR_ALIGN
        entries.append({"name": name, "shape": list(a.shape), "offset": off, "nbytes": int(a.nbytes),
                        "fnv1a64": f"0x{fnv1a64(a):016x}" if checksum else None})
        arrays.append((off, a))
        off += a.nbytes
    header = {"kind": kind, "meta": meta or {}, "tensors": entries}
    hj = json.dumps(header, sort_keys=True).encode()
    data_start = (16 + len(hj) + _DATA_ALIGN - 1) // _DATA_ALIGN * _DATA_ALIGN
    tmp = f"{path}.tmp.{os.getpid()}"
    with open(tmp, "wb") as f:
        f.write(MAGIC + struct.pack("<II", VERSION, len(hj)) + hj)
        for o, a in arrays:
            f.seek(data_start + o)
            f.write(a.tobytes())
        f.truncate(data_start + off)
        f.flush()
        os.fsync(f.fileno())
    os.replace(tmp, path)
    return header


def read_header(path: str) -> dict:
    with open(path, "rb") as f:
        head = f.read(16)
        if len(head) < 16 or head[:8] != MAGIC:
            raise ValueError(f"{path}: not a b200cnn weight file")
        ver, n = struct.unpack("<II", head[8:16])
        if ver != VERSION:
            raise ValueError(f"{path}: format version {ver}, this build reads {VERSION}")
        header = json.loads(f.read(n).decode())
    header["_data_start"] = (16 + n + _DATA_ALIGN - 1) // _DATA_ALIGN * _DATA_ALIGN
    return header


def load_weights(path: str, mmap: bool = False, verify: bool = True) -> Union[Blocks12Weights, "FullWeights"]:
    """Read a weight file.  ``mmap=True``: tensors are views over a private mapping of the file (no copy; a rank that slices
    its tensor-parallel shard only pages that shard in) and checksums are NOT verified unless ``verify`` is forced
    by reading everything anyway; ``mmap=False``: tensors own their memory and every checksum is checked."""
    header = read_header(path)
    base = header["_data_start"]
    size = os.path.getsize(path)
    got: Dict[str, torch.Tensor] = {}
    mm = np.memmap(path, dtype=np.uint8, mode="c") if mmap else None    # private copy-on-write mapping
    with open(path, "rb") as f:
        for e in header["tensors"]:
            lo, n = base + e["offset"], e["nbytes"]
            if lo + n > size:
                raise ValueError(f"{path}: tensor {e['name']} runs past the end of the file (truncated?)")
            count = 1
            for d in e["shape"]:
                count *= d
            if count * 4 != n:
                raise ValueError(f"{path}: tensor {e['name']} shape {e['shape']} does not match {n} bytes")
            if mmap:
                a = mm[lo:lo + n].view("<f4").reshape(e["shape"])
            else:
                f.seek(lo)
                a = np.frombuffer(f.read(n), dtype="<f4").reshape(e["shape"]).copy()
                if verify and e.get("fnv1a64"):
                    h = f"0x{fnv1a64(a):016x}"
                    if h != e["fnv1a64"]:
                        raise ValueError(f"{path}: checksum of {e['name']} is {h}, header says {e['fnv1a64']}")
            got[e["name"]] = torch.from_numpy(a)
    missing = [k for k in _B12 if k not in got]
    if missing:
        raise ValueError(f"{path}: missing tensors {missing}")
    b12 = Blocks12Weights(*(got[k] for k in _B12))
    if header["kind"] == "blocks12":
        return b12
    from ..models.alexnet_full import FullWeights
    missing = [k for k in _FULL if k not in got]
    if missing:
        raise ValueError(f"{path}: missing tensors {missing}")
    return FullWeights(b12, *(got[k] for k in _FULL))


def check_against(cfg, weights) -> None:
    """Shapes of a loaded file against the run's configuration (a wrong file must fail before any kernel runs)."""
    p1, p2 = cfg.conv1, cfg.conv2
    b12 = weights if isinstance(weights, Blocks12Weights) else weights.blocks12
    want = {"w1": (p1.K, cfg.C, p1.F, p1.F), "b1": (p1.K,), "w2": (p2.K, p1.K, p2.F, p2.F), "b2": (p2.K,)}
    for k, shp in want.items():
        have = tuple(getattr(b12, k).shape)
        if have != shp:
            raise ValueError(f"weight file: {k} has shape {have}, the configuration needs {shp}")


def resume_or_init(cfg, path: Optional[str], make, save_if_missing: bool = False, mmap: bool = False):
    """``make()`` synthesises the weights from the seed (the reference's behaviour).  With a ``path``: load it when it
    exists (and check it fits ``cfg``), else synthesise and, if asked, write it so the next start resumes from it.
    Returns ``(weights, "loaded"|"initialised"|"initialised+saved")``."""
    if path and os.path.exists(path):
        w = load_weights(path, mmap=mmap)
        check_against(cfg, w)
        return w, "loaded"
    w = make()
    if path and save_if_missing:
        save_weights(path, w, meta={"seed": int(getattr(cfg, "seed", 0)), "init": str(getattr(cfg, "init", ""))})
        return w, "initialised+saved"
    return w, "initialised"
