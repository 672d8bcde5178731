"""V1 — serial CPU (ref final_project/v1_serial/src/main.cpp, alexnet_serial.cpp:67-186).
Zero-argument run == reference: rand() init seeded with time(0), alpha/N LRN, whole image."""
from __future__ import annotations

import sys

from ..config import INIT_RAND, LRN_V1, config_from_args
from ..models.alexnet import forward_cpu
from ..models.init import make_problem
from ..ops.cpu import fnv1a64
from ..utils import logfmt as L
from .common import make_parser, WallTimer, weight_flags


def main(argv=None) -> int:
    ns = make_parser("v1_serial", init_default=INIT_RAND, lrn_default=LRN_V1).parse_args(argv)
    cfg = config_from_args(ns)
    L.emit(L.V1_BANNER)
    L.emit("Initializing data and parameters...")
    x, w = make_problem(cfg)
    w = weight_flags(cfg, ns, w)
    L.emit("Initialization complete.")
    L.emit("Starting AlexNet Serial Forward Pass...")

    def trace(layer, H, W, C):
        if layer in L.V1_LAYER_VERBS:
            pass
        L.emit(L.v1_dimensions(layer, H, W, C))
        nxt = {"Input": "After Conv1", "After Conv1": "After ReLU1", "After ReLU1": "After Pool1",
               "After Pool1": "After Conv2", "After Conv2": "After ReLU2",
               "After ReLU2": "After Pool2", "After Pool2": "After LRN2"}.get(layer)
        if nxt:
            L.emit(L.V1_LAYER_VERBS[nxt])

    with WallTimer() as t:
        y = forward_cpu(x, w, cfg, trace=trace)
        L.emit("Copying final result back to original buffer...")
    L.emit(L.v1_time(t.ms))
    flat = y.reshape(-1)
    L.emit(L.v1_values(flat[:10].tolist(), flat.numel()))
    L.emit("--- Serial execution finished ---")
    if not ns.quiet_extras:
        for line in L.extras(shape=y.shape[1:] if cfg.batch == 1 else y.shape, fnv=fnv1a64(y),
                             wall_ms=t.ms, seed=cfg.seed, batch=cfg.batch):
            L.emit(line)
    return 0


if __name__ == "__main__":
    sys.exit(main())
