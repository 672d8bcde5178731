"""Exact spatial (row) decomposition planner for conv->pool->conv->pool(->LRN) stacks.

The reference splits the *input* rows (``base = H/size``, first ``H % size`` ranks get
one more), exchanges ``F/2`` halo rows, and trims heuristically
(ref final_project/v2_mpi_only/2.2_scatter_halo/src/main.cpp:101-116,118-135,159-176;
final_project/v4_mpi_cuda/src/main_mpi_cuda.cpp:55-79,102-122).  That decomposition is
lossy: V4 returns 13 -> 8 -> 4 output rows at np 1/2/4.

This planner works backwards from the *output* instead.  Final rows are partitioned
with the same base+rem rule; every earlier layer's row range is derived through the
exact receptive-field arithmetic, so the union of all ranks' output is
exactly the np=1 output.  Two strategies:

* ``halo``      per-layer ownership: pool1 rows are disjointly owned, conv2 reads the
                rows it lacks from the owning peers (default; the fused peer-memory path);
* ``recompute`` overlapped tiling: each rank computes every pool1 row it needs itself from
                a larger input slab (the strategy of live V4, made exact).

Rows x batch.  13 final rows over 8 ranks split 2,2,2,2,2,1,1,1 (max/mean 1.23) and every rank then sits
next to two cuts.  ``row_groups=R`` (R divides the world) keeps the row decomposition inside groups of R
consecutive ranks and gives every group its own share of the batch (the batch dimension the reference
never had, SURVEY §2.4 "DP over batch"): 8 ranks as 2 row groups x 4 batch groups split 7/6 rows and a
rank only talks to the one neighbour that holds the other half of ITS images.  ``row_groups=None`` is
the reference's pure row split (R = world).

All intervals are half-open ``(lo, hi)`` with ``lo <= hi``; an empty range is ``(x, x)``.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Tuple

from ..config import NetConfig
from ..shapes import block12_shapes, Block12Shapes

Range = Tuple[int, int]


def split_rows(n: int, parts: int) -> List[Range]:
    """``base = n/parts`` rows each, the first ``n % parts`` ranks get one extra
    (ref v2.2 main.cpp:102-109).  Ranks beyond ``n`` get empty ranges."""
    base, rem = divmod(n, parts)
    out, off = [], 0
    for r in range(parts):
        k = base + (1 if r < rem else 0)
        out.append((off, off + k))
        off += k
    return out


def window_in_rows(out: Range, F: int, S: int, P: int, D_in: int) -> Range:
    """Input rows touched by output rows ``out`` of a window op (conv or pool),
    clipped to the tensor (rows outside are the op's own zero padding)."""
    lo, hi = out
    if hi <= lo:
        return (0, 0)
    a = lo * S - P
    b = (hi - 1) * S - P + F          # exclusive
    return (max(a, 0), min(b, D_in))


def _intersect(a: Range, b: Range) -> Range:
    lo, hi = max(a[0], b[0]), min(a[1], b[1])
    return (lo, hi) if hi > lo else (lo, lo)


def _len(r: Range) -> int:
    return max(0, r[1] - r[0])


@dataclass
class RankPlan:
    rank: int
    world: int
    out_rows: Range          # owned final rows (pool2 / LRN2)
    conv2_rows: Range        # conv2 rows this rank computes
    pool1_need: Range        # pool1 rows conv2 reads (clipped; outside = zero pad)
    pool1_own: Range         # pool1 rows this rank produces
    conv1_rows: Range        # conv1 rows computed for pool1_own
    in_rows: Range           # input rows conv1 reads (the "scatter" slab)
    halo: List[Tuple[int, int, int]] = field(default_factory=list)    # (peer, lo, hi) pool1 rows read from peer
    serves: List[Tuple[int, int, int]] = field(default_factory=list)  # (peer, lo, hi) pool1 rows peer reads from me
    images: Range = (0, 0)   # images [n0, n1) of the global batch this rank works on (its batch group's share)
    row_group: int = 0       # position inside the group of ranks that split the rows of the same images
    batch_group: int = 0

    @property
    def active(self) -> bool:
        return _len(self.out_rows) > 0 and _len(self.images) > 0

    @property
    def n_images(self) -> int:
        return _len(self.images)


@dataclass
class Plan:
    cfg: NetConfig
    shapes: Block12Shapes
    world: int
    strategy: str
    ranks: List[RankPlan]
    row_groups: int = 0      # R: ranks that split the rows of one image set (== world for the pure row split)
    batch_groups: int = 1    # B: world == R * B, rank == batch_group * R + row_group

    def owner_of_pool1_row(self, p: int, batch_group: int = 0) -> int:
        for rp in self.ranks:
            if rp.batch_group == batch_group and rp.pool1_own[0] <= p < rp.pool1_own[1]:
                return rp.rank
        raise IndexError(p)

    def summary(self) -> str:
        lines = [f"plan world={self.world} = {self.row_groups} row groups x {self.batch_groups} batch groups, "
                 f"strategy={self.strategy} H={self.shapes.H} -> out {self.shapes.Hp2}x{self.shapes.Wp2}x{self.shapes.K2}"]
        for r in self.ranks:
            lines.append(f"  r{r.rank}: images{r.images} out{r.out_rows} conv2{r.conv2_rows} need{r.pool1_need} "
                         f"own{r.pool1_own} conv1{r.conv1_rows} in{r.in_rows} halo{r.halo}")
        return "\n".join(lines)


def default_row_groups(world: int) -> int:
    """Rows x batch shape used when the caller asks for ``row_groups="auto"``: two row groups whenever the
    world is even (every rank has exactly one halo neighbour and the 13 final rows of AlexNet split 7/6),
    otherwise the pure row split."""
    return 2 if world >= 2 and world % 2 == 0 else world


def make_plan(cfg: NetConfig, world: int, strategy: str = "halo", row_groups=None) -> Plan:
    """``row_groups``: None / 0 = pure row split over all ranks (the reference's decomposition),
    "auto" = :func:`default_row_groups`, or an explicit divisor R of ``world``."""
    if strategy not in ("halo", "recompute"):
        raise ValueError(strategy)
    if row_groups == "auto":
        row_groups = default_row_groups(world)
    R = int(row_groups) if row_groups else world
    if R < 1 or world % R != 0:
        raise ValueError(f"row_groups={R} does not divide world={world}")
    B = world // R
    group = _make_row_plan(cfg, R, strategy)
    shares = split_rows(cfg.batch, B)
    ranks: List[RankPlan] = []
    for bg in range(B):
        for q in group:
            ranks.append(RankPlan(
                rank=bg * R + q.rank, world=world, out_rows=q.out_rows, conv2_rows=q.conv2_rows,
                pool1_need=q.pool1_need, pool1_own=q.pool1_own, conv1_rows=q.conv1_rows, in_rows=q.in_rows,
                halo=[(bg * R + p, lo, hi) for p, lo, hi in q.halo],
                serves=[(bg * R + p, lo, hi) for p, lo, hi in q.serves],
                images=shares[bg], row_group=q.rank, batch_group=bg))
    return Plan(cfg, block12_shapes(cfg), world, strategy, ranks, R, B)


def _make_row_plan(cfg: NetConfig, world: int, strategy: str) -> List[RankPlan]:
    """The row decomposition of one image set over ``world`` ranks (ranks numbered inside the group)."""
    sh = block12_shapes(cfg)
    p1, p2 = cfg.conv1, cfg.conv2
    finals = split_rows(sh.Hp2, world)

    # pool1 ownership: pool1 row p feeds conv2 row p (stride-1 centre) which falls in
    # pool2 row floor(p*S2 / S_pool2) (clipped) -> owner of that final row.  Monotone, so
    # every rank's owned rows are contiguous and ranks without output own nothing.
    def final_row_of_pool1(p: int) -> int:
        c = min(max((p + p2.P - p2.F // 2) // max(p2.S, 1), 0), sh.Hc2 - 1) if sh.Hc2 else 0
        return min(c // p2.S_pool, sh.Hp2 - 1)

    own = [[None, None] for _ in range(world)]
    if sh.Hp2 > 0:
        row_owner = {}
        for r, (a, b) in enumerate(finals):
            for f in range(a, b):
                row_owner[f] = r
        for p in range(sh.Hp1):
            r = row_owner[final_row_of_pool1(p)]
            if own[r][0] is None:
                own[r][0] = p
            own[r][1] = p + 1

    ranks: List[RankPlan] = []
    for r, out in enumerate(finals):
        conv2 = window_in_rows(out, p2.F_pool, p2.S_pool, 0, sh.Hc2)
        need = window_in_rows(conv2, p2.F, p2.S, p2.P, sh.Hp1)
        if strategy == "halo":
            o = (own[r][0], own[r][1]) if own[r][0] is not None else (need[0], need[0])
        else:
            o = need
        conv1 = window_in_rows(o, p1.F_pool, p1.S_pool, 0, sh.Hc1)
        inr = window_in_rows(conv1, p1.F, p1.S, p1.P, sh.H)
        ranks.append(RankPlan(r, world, out, conv2, need, o, conv1, inr))

    if strategy == "halo":
        for rp in ranks:
            if _len(rp.out_rows) == 0:
                continue
            for other in ranks:
                if other.rank == rp.rank or _len(other.pool1_own) == 0:
                    continue
                ov = _intersect(rp.pool1_need, other.pool1_own)
                if _len(ov):
                    rp.halo.append((other.rank, ov[0], ov[1]))
                    other.serves.append((rp.rank, ov[0], ov[1]))
    return ranks


def validate_plan(plan: Plan) -> None:
    """Invariants (also exercised by hypothesis tests): output rows tile [0,Hp2) exactly,
    pool1 needs are covered by own+halo, halo rows come from their owner."""
    sh = plan.shapes
    assert plan.row_groups * plan.batch_groups == plan.world == len(plan.ranks)
    imgs = []
    for bg in range(plan.batch_groups):
        members = [rp for rp in plan.ranks if rp.batch_group == bg]
        assert [rp.rank for rp in members] == list(range(bg * plan.row_groups, (bg + 1) * plan.row_groups))
        assert len({rp.images for rp in members}) == 1, "a row group must share one image range"
        imgs.extend(range(*members[0].images))
        covered = []
        for rp in members:
            covered.extend(range(*rp.out_rows))
        assert covered == list(range(sh.Hp2)), "final rows are not an exact partition"
        if plan.strategy == "halo":
            owned = []
            for rp in members:
                owned.extend(range(*rp.pool1_own))
            used = sorted(set(p for rp in members for p in range(*rp.pool1_need)))
            assert owned == sorted(set(owned)), "pool1 ownership overlaps"
            assert set(used) <= set(owned), "a needed pool1 row has no owner"
    assert imgs == list(range(plan.cfg.batch)), "batch groups are not an exact partition of the images"
    for rp in plan.ranks:
        if _len(rp.out_rows) == 0:
            continue
        have = set(range(*rp.pool1_own))
        for (peer, lo, hi) in rp.halo:
            q = plan.ranks[peer]
            assert q.batch_group == rp.batch_group and q.images == rp.images, "halo peer works on other images"
            po = q.pool1_own
            assert po[0] <= lo and hi <= po[1], "halo rows not owned by the named peer"
            have |= set(range(lo, hi))
        assert set(range(*rp.pool1_need)) <= have, "conv2 input rows not covered"
