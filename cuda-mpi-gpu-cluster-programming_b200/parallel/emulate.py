"""Single-process emulation of the row-decomposed forward on CPU: every rank's work is run
in turn and halo rows are copied between the per-rank slabs.  Used by tests (np in 1..8
without spawning processes) and as the model for the multi-process CPU runner."""
from __future__ import annotations

import torch

from ..config import NetConfig
from ..models.alexnet import block1_rank_cpu, block2_rank_cpu
from ..models.init import Blocks12Weights
from .planner import make_plan, Plan


def assemble_pool1(plan: Plan, rank: int, own_slabs) -> torch.Tensor:
    """Rows ``pool1_need`` of rank ``rank`` = its own rows + halo rows from their owners."""
    rp = plan.ranks[rank]
    n0, n1 = rp.pool1_need
    N = own_slabs[rank].shape[0]          # the images of this rank's batch group (halo peers hold the same ones)
    sh = plan.shapes
    buf = torch.zeros((N, n1 - n0, sh.Wp1, sh.K1), dtype=torch.float32)
    o0, o1 = rp.pool1_own
    lo, hi = max(o0, n0), min(o1, n1)
    if hi > lo:
        buf[:, lo - n0:hi - n0] = own_slabs[rank][:, lo - o0:hi - o0]
    for peer, lo, hi in rp.halo:
        pb = plan.ranks[peer].pool1_own[0]
        buf[:, lo - n0:hi - n0] = own_slabs[peer][:, lo - pb:hi - pb]
    return buf


def forward_cpu_decomposed(x: torch.Tensor, w: Blocks12Weights, cfg: NetConfig, world: int,
                           strategy: str = "halo", row_groups=None) -> torch.Tensor:
    plan = make_plan(cfg.with_(batch=x.shape[0], H=x.shape[1], W=x.shape[2]), world, strategy, row_groups)
    sh = plan.shapes
    own = []
    for rp in plan.ranks:
        i0, i1 = rp.in_rows
        a, b = rp.images
        if b <= a:                       # more batch groups than images: this group has nothing to do
            own.append(None)
            continue
        own.append(block1_rank_cpu(x[a:b, i0:i1].contiguous(), i0, w, cfg, rp, sh.H))
    out = torch.empty((x.shape[0], sh.Hp2, sh.Wp2, sh.K2), dtype=torch.float32)
    for rp in plan.ranks:
        if not rp.active:
            continue
        slab = assemble_pool1(plan, rp.rank, own)
        y = block2_rank_cpu(slab, rp.pool1_need[0], w, cfg, rp, sh.Hp1)
        out[rp.images[0]:rp.images[1], rp.out_rows[0]:rp.out_rows[1]] = y
    return out
