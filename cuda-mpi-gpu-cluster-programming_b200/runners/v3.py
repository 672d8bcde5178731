"""V3 — single GPU (ref final_project/v3_cuda_only/src/main_cuda.cpp:12-44).

Zero-argument run == reference: constant init, raw-alpha LRN (V3/V4 convention),
one 227x227x3 image, wall clock around the whole forward call (the reference's timed region
contains context creation, allocation and the pageable copies, ref alexnet_cuda.cu:54-94; ours
contains session creation, H2D from pinned memory, two fused kernels and D2H).
``--iters K`` adds warm, device-timed repetitions as key=value lines."""
from __future__ import annotations

import sys

import torch

from ..config import INIT_CONST, LRN_V3, config_from_args
from ..models.init import make_problem
from ..models.session import Session
from ..ops.cpu import fnv1a64
from ..utils import logfmt as L
from .common import make_parser, WallTimer, fail, weight_flags


def warm_device_ms(sess: Session, iters: int) -> float:
    e = sess.engine
    for _ in range(3):
        e.forward_enqueue(0, 0)
    torch.cuda.synchronize(e.device)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(iters):
        e.forward_enqueue(0, 0)
    ev1.record()
    ev1.synchronize()
    return ev0.elapsed_time(ev1) / iters


def main(argv=None) -> int:
    ns = make_parser("v3_cuda_only", init_default=INIT_CONST, lrn_default=LRN_V3).parse_args(argv)
    cfg = config_from_args(ns, seed_default=1)
    if not torch.cuda.is_available():
        fail(0, "CUDA no CUDA-capable device is detected")
    x, w = make_problem(cfg)
    w = weight_flags(cfg, ns, w)
    xp = x.pin_memory()
    with WallTimer() as t:
        sess = Session(cfg, w, device="cuda:0", precision=cfg.precision)
        y = sess.infer(xp)
    flat = y.reshape(-1)
    for line in L.v3_lines(flat[:10].tolist(), t.ms):
        L.emit(line)
    if not ns.quiet_extras:
        kv = {}
        if ns.iters > 0:
            ms = warm_device_ms(sess, ns.iters)
            kv = dict(device_forward_ms_warm=ms, images_per_s_warm=cfg.batch / (ms / 1e3))
        for line in L.extras(shape=y.shape[1:] if cfg.batch == 1 else y.shape, fnv=fnv1a64(y),
                             wall_ms=t.ms, precision=cfg.precision, batch=cfg.batch, **kv):
            L.emit(line)
    return 0


if __name__ == "__main__":
    sys.exit(main())
