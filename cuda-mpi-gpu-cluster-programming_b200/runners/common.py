"""Shared pieces of the version runners (the L4 entry points)."""
from __future__ import annotations

import argparse
import sys
import time

from ..config import add_cli_flags, INIT_CONST, LRN_V1


def make_parser(prog: str, *, init_default: str = INIT_CONST, lrn_default: str = LRN_V1):
    ap = argparse.ArgumentParser(
        prog=prog, description="With no arguments this reproduces the reference binary "
        "(which ignores argv; SURVEY §0.2).  All flags are additions.")
    add_cli_flags(ap, init_default=init_default, lrn_default=lrn_default)
    ap.add_argument("--strategy", choices=["halo", "recompute"], default="halo")
    ap.add_argument("--scatter", choices=["p2p", "sharded"], default="p2p")
    ap.add_argument("--row-groups", default="rows",
                    help="rows (default: the reference's pure row split), auto (2 row groups x np/2 batch groups) "
                         "or a divisor R of np: R ranks split the rows of one share of the batch")
    ap.add_argument("--gather", choices=["rank0", "host"], default="rank0",
                    help="rank0: rows are stored into rank 0's device tensor over NVLink; host: every rank copies "
                         "its rows into one shared pinned host tensor over its own PCIe link")
    ap.add_argument("--quiet-extras", action="store_true", help="print only the reference's lines")
    return ap


def weight_flags(cfg, ns, w, rank: int = 0):
    """--load-weights / --save-weights for blocks 1-2: returns the Blocks12Weights the run uses.  A file of the
    full-network kind is accepted too (its first two blocks are taken)."""
    from ..utils import checkpoint as ck
    load = getattr(ns, "load_weights", None)
    if load:
        try:
            got = ck.load_weights(load)
            ck.check_against(cfg, got)
        except (OSError, ValueError) as ex:
            fail(rank, f"--load-weights {load}: {ex}")
        w = got if not hasattr(got, "blocks12") else got.blocks12
    save = getattr(ns, "save_weights", None)
    if save and rank == 0:
        ck.save_weights(save, w, meta={"seed": int(cfg.seed), "init": str(cfg.init), "kind": "blocks12"})
    return w


class WallTimer:
    def __enter__(self):
        self.t0 = time.perf_counter()
        return self

    def __exit__(self, *a):
        self.ms = (time.perf_counter() - self.t0) * 1e3


def fail(rank: int, msg: str, code: int = 1):
    sys.stderr.write(f"[Rank {rank}] {msg}\n")
    sys.stderr.flush()
    raise SystemExit(code)
