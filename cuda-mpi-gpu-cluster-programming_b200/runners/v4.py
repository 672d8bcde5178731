"""V4 / V5 — row-decomposed forward across GPUs, one process per GPU.

V4 in the reference = MPI + CUDA with host staging (ref v4_mpi_cuda/src/main_mpi_cuda.cpp:20-163);
V5 = the CUDA-aware variant that was specified (README.md:158-166) but never written.
Both map onto the same engine here, because the host staging simply does not exist any more:

* ``v4`` keeps the reference's data placement: rank 0 owns the image; the other ranks pull their
  rows out of rank 0's device buffer from inside conv1 (scatter="p2p").
* ``v5`` lets every rank feed its own row slab from its own pinned host memory (scatter="sharded"),
  which is what "device pointers end to end" buys on an 8-GPU node: eight PCIe links instead of one.

Unlike the reference (13 -> 8 -> 4 output rows at np 1/2/4) the decomposition is exact.  Prints the
V4 lines: ``Final Output Shape:``, ``Final Output (first 10 values):``, ``AlexNet MPI+CUDA Forward
Pass completed in X ms`` (timed like the reference: barrier .. barrier around the whole run).
"""
from __future__ import annotations

import sys
import time

import torch

from ..config import INIT_CONST, LRN_V3, config_from_args
from ..models.init import make_problem
from ..models.session import Session
from ..ops.cpu import fnv1a64
from ..parallel import dist as bdist
from ..utils import logfmt as L
from .common import make_parser, fail, weight_flags
from .v3 import warm_device_ms


def main(argv=None, variant: str = "v4") -> int:
    ap = make_parser(f"{variant}_mpi_cuda", init_default=INIT_CONST, lrn_default=LRN_V3)
    ns = ap.parse_args(argv)
    cfg = config_from_args(ns, seed_default=1)
    if not torch.cuda.is_available():
        fail(bdist.env_rank_world()[0], "CUDA no CUDA-capable device is detected")
    scatter = "sharded" if variant == "v5" else ns.scatter
    world_env = bdist.env_rank_world()[1]
    if world_env > torch.cuda.device_count():
        return _oversubscribed(ns, cfg, variant, scatter)
    ctx = bdist.init()
    try:
        x, w = make_problem(cfg)                  # identical on every rank (seeded): no weight bcast
        w = weight_flags(cfg, ns, w, ctx.rank)    # or the same file read by every rank
        ctx.barrier()
        t0 = time.perf_counter()
        sess = Session(cfg, w, dist=ctx if ctx.world > 1 else None, device=ctx.device,
                       precision=cfg.precision, strategy=ns.strategy, scatter=scatter,
                       row_groups=_row_groups(ns), gather=ns.gather)
        xin = sess.my_input(x)
        xin = xin.contiguous().pin_memory() if xin is not None else None
        y = sess.infer(xin)
        ctx.barrier()
        ms = (time.perf_counter() - t0) * 1e3
        kv = {}
        if ns.iters > 0:
            dms = ctx.max_float(warm_device_ms(sess, ns.iters))
            kv = dict(device_forward_ms_warm=dms, images_per_s_warm=cfg.batch / (dms / 1e3))
    except SystemExit:
        raise
    except Exception as ex:                       # CUDA_CHECK -> MPI_Abort semantics
        L.rank_err(ctx.rank, f"CUDA {type(ex).__name__}: {ex}")
        ctx.abort(1)
    if ctx.rank == 0:
        flat = y.reshape(-1)
        rows = y.shape[1] if cfg.batch == 1 else y.shape[0] * y.shape[1]
        for line in L.v4_shape_values(rows, y.shape[2], y.shape[3], flat[:10].tolist()):
            L.emit(line)
        L.emit(L.v4_time(ms, "MPI+CUDA" if variant == "v4" else "MPI+CUDA-Aware"))
        if not ns.quiet_extras:
            for line in L.extras(fnv=fnv1a64(y), wall_ms=ms, np=ctx.world, precision=cfg.precision,
                                 scatter=scatter, strategy=ns.strategy, batch=cfg.batch, **kv):
                L.emit(line)
    ctx.finalize()
    return 0


def _row_groups(ns):
    rg = getattr(ns, "row_groups", None)
    if rg in (None, "", "rows"):
        return None                      # the reference's decomposition: rows only
    return "auto" if rg == "auto" else int(rg)


def _oversubscribed(ns, cfg, variant, scatter) -> int:
    """More ranks than GPUs (the reference happily runs `mpirun -np 4` on one GPU, ref
    scripts/common_test_utils.sh:275).  Ranks that spin on each other's flags must never share a
    GPU (B200_PROFILING.md: Xid 109), so rank 0 runs all ranks' kernels itself, layer by layer,
    through the single-device emulation of the exact same protocol; the other processes only take
    part in the barriers.  Output and log lines are identical to a real multi-GPU run."""
    from ..models.engine import EmulatedWorld
    ctx = bdist.init(device="cpu")
    if ctx.rank == 0:
        # same footing as the one-GPU-per-rank path, where bdist.init() selects the device (CUDA 12 creates the
        # primary context there) and the data is generated before the timed region opens
        torch.cuda.set_device(0)
        from ..ops import cuda as cu
        cu.lib()
        x, w = make_problem(cfg)
        w = weight_flags(cfg, ns, w, 0)
    ctx.barrier()
    t0 = time.perf_counter()
    y = None
    if ctx.rank == 0:
        ew = EmulatedWorld(cfg, w, ctx.world, "cuda:0", precision=cfg.precision, strategy=ns.strategy,
                           scatter=scatter, row_groups=_row_groups(ns))
        y = ew.forward(x.pin_memory().to("cuda:0", non_blocking=True)).cpu()
    ctx.barrier()
    ms = (time.perf_counter() - t0) * 1e3
    if ctx.rank == 0:
        flat = y.reshape(-1)
        rows = y.shape[1] if cfg.batch == 1 else y.shape[0] * y.shape[1]
        for line in L.v4_shape_values(rows, y.shape[2], y.shape[3], flat[:10].tolist()):
            L.emit(line)
        L.emit(L.v4_time(ms, "MPI+CUDA" if variant == "v4" else "MPI+CUDA-Aware"))
        if not ns.quiet_extras:
            for line in L.extras(fnv=fnv1a64(y), wall_ms=ms, np=ctx.world, precision=cfg.precision,
                                 scatter=scatter, strategy=ns.strategy, batch=cfg.batch,
                                 mode="ranks-emulated-on-one-gpu"):
                L.emit(line)
    ctx.finalize()
    return 0


def main_v5(argv=None) -> int:
    return main(argv, "v5")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] in ("v4", "v5"):
        sys.exit(main(sys.argv[2:], sys.argv[1]))
    sys.exit(main(sys.argv[1:]))
