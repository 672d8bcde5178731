"""Typed configuration: per-block hyper-parameters and run options.

Defaults equal the constants every reference ``main`` hard-codes
(ref: final_project/v1_serial/src/main.cpp:18-43,
final_project/v4_mpi_cuda/src/main_mpi_cuda.cpp:146-150).  The reference has no
runtime flags at all; everything configurable here is an addition whose
zero-argument default reproduces the reference.
"""
from __future__ import annotations

import argparse
from dataclasses import dataclass, asdict, replace
from typing import Optional


@dataclass(frozen=True)
class LayerParams:
    """One conv block: conv (K filters FxF, stride S, pad P) -> ReLU -> max-pool
    (F_pool, S_pool) -> optional cross-channel LRN (N_lrn, alpha, beta, k_lrn).

    Field names follow ref final_project/v1_serial/include/alexnet.hpp:9-24.  Weights
    and biases are *not* stored here (the reference keeps host vectors inside the
    struct); tensors live in :class:`b200cnn.models.alexnet.Blocks12Weights`.
    """

    K: int
    F: int
    S: int
    P: int
    F_pool: int = 3
    S_pool: int = 2
    N_lrn: int = 0          # 0 => block has no LRN
    alpha: float = 1e-4
    beta: float = 0.75
    k_lrn: float = 2.0


CONV1 = LayerParams(K=96, F=11, S=4, P=0, F_pool=3, S_pool=2)
CONV2 = LayerParams(K=256, F=5, S=1, P=2, F_pool=3, S_pool=2,
                    N_lrn=5, alpha=1e-4, beta=0.75, k_lrn=2.0)

# LRN conventions: V1/V2 divide alpha by the window size, V3/V4 do not.
LRN_V1 = "alpha_over_n"     # canonical: V1 is the oracle
LRN_V3 = "alpha_raw"

# Initialisers
INIT_CONST = "const"        # V2-V4: input 1.0, weights 0.01, biases 0.0
INIT_RAND = "rand"          # V1: glibc rand() stream, biases 0.1


@dataclass(frozen=True)
class NetConfig:
    H: int = 227
    W: int = 227
    C: int = 3
    batch: int = 1
    conv1: LayerParams = CONV1
    conv2: LayerParams = CONV2
    lrn_mode: str = LRN_V1
    init: str = INIT_CONST
    seed: int = 1
    # numeric path on GPU: "simt" exact FP32, "tf32" tcgen05 single pass
    precision: str = "tf32"

    def with_(self, **kw) -> "NetConfig":
        return replace(self, **kw)

    def to_dict(self) -> dict:
        return asdict(self)


def default_config(**kw) -> NetConfig:
    return NetConfig(**kw)


def add_cli_flags(ap: argparse.ArgumentParser, *, init_default: str = INIT_CONST,
                  lrn_default: str = LRN_V1) -> None:
    """Flags shared by every version runner.  All optional: with none given the runner
    behaves like the reference binary (which ignores argv)."""
    ap.add_argument("--height", type=int, default=227)
    ap.add_argument("--width", type=int, default=227)
    ap.add_argument("--channels", type=int, default=3)
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--init", choices=[INIT_CONST, INIT_RAND], default=init_default)
    ap.add_argument("--seed", type=int, default=None,
                    help="seed for --init rand (default: time(0), like V1)")
    ap.add_argument("--lrn", choices=[LRN_V1, LRN_V3], default=lrn_default)
    ap.add_argument("--precision", choices=["simt", "tf32"], default="tf32")
    ap.add_argument("--iters", type=int, default=0,
                    help="extra warm device-timed iterations (adds key=value lines)")
    ap.add_argument("--load-weights", default=None, metavar="FILE",
                    help="take the parameters from a weight file (utils/checkpoint.py) instead of --init")
    ap.add_argument("--save-weights", default=None, metavar="FILE",
                    help="write the parameters this run used to a weight file (rank 0)")


def config_from_args(ns: argparse.Namespace, seed_default: Optional[int] = None) -> NetConfig:
    import time
    seed = ns.seed
    if seed is None:
        seed = seed_default if seed_default is not None else int(time.time())
    return NetConfig(H=ns.height, W=ns.width, C=ns.channels, batch=ns.batch,
                     lrn_mode=ns.lrn, init=ns.init, seed=seed, precision=ns.precision)
