"""V2.1 (broadcast-all) and V2.2 (scatter + halo) — CPU, one process per rank over gloo.

* V2.1: rank 0 builds the data, everything is broadcast, EVERY rank computes the whole image,
  rank 0 prints (ref v2_mpi_only/2.1_broadcast_all/src/main.cpp:10-109; there is no gather).
* V2.2: the spatial-decomposition engine (ref v2_mpi_only/2.2_scatter_halo/src/main.cpp:99-251):
  scatter input rows, Conv1/ReLU/Pool1 locally, exchange pool1 halo rows with neighbours,
  Conv2/ReLU/Pool2/LRN2, gather on rank 0.  The reference's version is lossy (14 rows at np2,
  10 at np4); this one uses the exact planner, so np>1 output == np1 output.

Timed region as in the reference: from after the broadcasts (barrier) to after the gather
(barrier); prints ``shape:`` / ``Sample values:`` / ``Execution Time:``.
"""
from __future__ import annotations

import sys
import time

import torch

from ..config import INIT_CONST, LRN_V1, config_from_args
from ..models.alexnet import forward_cpu, block1_rank_cpu, block2_rank_cpu
from ..models.init import make_problem, Blocks12Weights
from ..ops.cpu import fnv1a64
from ..parallel import dist as bdist
from ..parallel.planner import make_plan
from ..shapes import block12_shapes
from ..utils import logfmt as L
from .common import make_parser, weight_flags

TAG_SCATTER, TAG_HALO, TAG_GATHER = 10, 2, 20     # ref uses tags 0/1 (halo1) and 2/3 (halo2)


def _bcast_problem(ctx, cfg, need_input: bool, ns=None):
    """rank 0 initialises (or loads --load-weights), everybody receives weights (and the input for V2.1)."""
    p1, p2 = cfg.conv1, cfg.conv2
    if ctx.rank == 0:
        x, w = make_problem(cfg)
        if ns is not None:
            w = weight_flags(cfg, ns, w, 0)
    else:
        x = torch.empty((cfg.batch, cfg.H, cfg.W, cfg.C)) if need_input else None
        w = Blocks12Weights(torch.empty((p1.K, cfg.C, p1.F, p1.F)), torch.empty(p1.K),
                            torch.empty((p2.K, p1.K, p2.F, p2.F)), torch.empty(p2.K))
    for t in (w.w1, w.b1, w.w2, w.b2):
        ctx.bcast(t, 0)
    if need_input:
        ctx.bcast(x, 0)
    return x, w


def run_v21(ctx, cfg, ns=None):
    x, w = _bcast_problem(ctx, cfg, need_input=True, ns=ns)
    ctx.barrier()
    t0 = time.perf_counter()
    y = forward_cpu(x, w, cfg)
    ctx.barrier()
    return y if ctx.rank == 0 else None, (time.perf_counter() - t0) * 1e3


def run_v22(ctx, cfg, strategy="halo", ns=None):
    x, w = _bcast_problem(ctx, cfg, need_input=False, ns=ns)
    sh = block12_shapes(cfg)
    plan = make_plan(cfg, ctx.world, strategy)
    rp = plan.ranks[ctx.rank]
    ctx.barrier()
    t0 = time.perf_counter()
    if ctx.world == 1:
        y = forward_cpu(x, w, cfg)
        return y, (time.perf_counter() - t0) * 1e3

    # -- scatter: every rank gets exactly the input rows its conv1 rows need (MPI_Scatterv) ----
    i0, i1 = rp.in_rows
    if ctx.rank == 0:
        reqs = [ctx.isend(x[:, q.in_rows[0]:q.in_rows[1]].contiguous(), q.rank, TAG_SCATTER)
                for q in plan.ranks[1:] if q.in_rows[1] > q.in_rows[0]]
        mine = x[:, i0:i1].contiguous()
        for r in reqs:
            r.wait()
    else:
        mine = torch.empty((cfg.batch, i1 - i0, cfg.W, cfg.C))
        if i1 > i0:
            ctx.recv(mine, 0, TAG_SCATTER)

    # -- block 1 on the owned pool1 rows --------------------------------------------------------
    own = block1_rank_cpu(mine, i0, w, cfg, rp, sh.H)

    # -- halo exchange of pool1 rows (post receives, then sends, then wait: PROBLEMS.txt:47-48) --
    n0, n1 = rp.pool1_need
    slab = torch.zeros((cfg.batch, n1 - n0, sh.Wp1, sh.K1))
    o0, o1 = rp.pool1_own
    lo, hi = max(o0, n0), min(o1, n1)
    if hi > lo:
        slab[:, lo - n0:hi - n0] = own[:, lo - o0:hi - o0]
    rbufs, reqs = [], []
    for peer, a, b in rp.halo:
        buf = torch.empty((cfg.batch, b - a, sh.Wp1, sh.K1))
        rbufs.append((a, b, buf))
        reqs.append(ctx.irecv(buf, peer, TAG_HALO))
    for peer, a, b in rp.serves:
        reqs.append(ctx.isend(own[:, a - o0:b - o0].contiguous(), peer, TAG_HALO))
    for r in reqs:
        r.wait()
    for a, b, buf in rbufs:
        slab[:, a - n0:b - n0] = buf

    # -- block 2 on the owned final rows, gather on rank 0 (MPI_Gatherv) ------------------------
    y_loc = block2_rank_cpu(slab, n0, w, cfg, rp, sh.Hp1) if rp.active else None
    y = None
    if ctx.rank == 0:
        y = torch.empty((cfg.batch, sh.Hp2, sh.Wp2, sh.K2))
        if rp.active:
            y[:, rp.out_rows[0]:rp.out_rows[1]] = y_loc
        for q in plan.ranks[1:]:
            if q.active:
                buf = torch.empty((cfg.batch, q.out_rows[1] - q.out_rows[0], sh.Wp2, sh.K2))
                ctx.recv(buf, q.rank, TAG_GATHER)
                y[:, q.out_rows[0]:q.out_rows[1]] = buf
    elif rp.active:
        ctx.send(y_loc, 0, TAG_GATHER)
    ctx.barrier()
    return y, (time.perf_counter() - t0) * 1e3


def main(argv=None, variant="2.2") -> int:
    ap = make_parser(f"v2_{variant}", init_default=INIT_CONST, lrn_default=LRN_V1)
    ns = ap.parse_args(argv)
    cfg = config_from_args(ns, seed_default=1)
    ctx = bdist.init(device="cpu")
    try:
        if variant == "2.1":
            y, ms = run_v21(ctx, cfg, ns=ns)
        else:
            y, ms = run_v22(ctx, cfg, ns.strategy, ns=ns)
    except Exception as ex:            # fail-stop: MPI_Abort semantics
        L.rank_err(ctx.rank, f"{type(ex).__name__}: {ex}")
        ctx.abort(1)
    if ctx.rank == 0:
        sh = block12_shapes(cfg)
        if y.numel() != cfg.batch * sh.out_numel:
            L.emit(L.v22_size_warning(y.numel(), *sh.out_hwc), file=sys.stderr)
        flat = y.reshape(-1)
        for line in L.v2_lines(sh.Hp2, sh.Wp2, sh.K2, flat[:5].tolist(), ms):
            L.emit(line)
        if not ns.quiet_extras:
            for line in L.extras(fnv=fnv1a64(y), wall_ms=ms, np=ctx.world, batch=cfg.batch):
                L.emit(line)
    ctx.finalize()
    return 0


def main_v21(argv=None) -> int:
    return main(argv, "2.1")


def main_v22(argv=None) -> int:
    return main(argv, "2.2")


if __name__ == "__main__":
    sys.exit(main(sys.argv[2:], sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1] in ("2.1", "2.2")
             else main(sys.argv[1:]))
