"""Planner properties (hypothesis) and exactness of the decomposition: np in {1..8} output ==
np=1 output — the property the reference's V2.2/V4 decomposition fails."""
import pytest
import torch
from hypothesis import given, settings, strategies as st

from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem
from b200cnn.models.alexnet import forward_cpu
from b200cnn.parallel.planner import make_plan, validate_plan, split_rows
from b200cnn.parallel.emulate import forward_cpu_decomposed


def test_split_rows_matches_reference_rule():
    # base = H/size, first H%size ranks get one more (ref v2.2 main.cpp:102-109)
    assert split_rows(227, 2) == [(0, 114), (114, 227)]
    assert [b - a for a, b in split_rows(13, 8)] == [2, 2, 2, 2, 2, 1, 1, 1]
    assert [b - a for a, b in split_rows(3, 5)] == [1, 1, 1, 0, 0]


def test_plan_227_np2_rows():
    plan = make_plan(default_config(), 2)
    validate_plan(plan)
    r0, r1 = plan.ranks
    assert r0.out_rows == (0, 7) and r1.out_rows == (7, 13)
    assert r0.conv2_rows == (0, 15) and r1.conv2_rows == (14, 27)
    assert r0.pool1_need == (0, 17) and r1.pool1_need == (12, 27)
    assert r0.pool1_own == (0, 14) and r1.pool1_own == (14, 27)
    assert r0.halo == [(1, 14, 17)] and r1.halo == [(0, 12, 14)]
    # receptive field of App. A.3: input rows [16a-16, 16b+50] clipped (recompute strategy)
    rc = make_plan(default_config(), 2, "recompute")
    assert rc.ranks[1].in_rows == (16 * 7 - 16, 227)


@settings(max_examples=150, deadline=None)
@given(H=st.integers(35, 700), world=st.integers(1, 12),
       strategy=st.sampled_from(["halo", "recompute"]))
def test_plan_invariants(H, world, strategy):
    plan = make_plan(default_config(H=H, W=67), world, strategy)
    validate_plan(plan)
    for rp in plan.ranks:
        if rp.active:
            # input slab is large enough for every conv1 row the rank computes
            q0, q1 = rp.conv1_rows
            assert rp.in_rows[0] <= 4 * q0 and 4 * (q1 - 1) + 11 <= rp.in_rows[1]


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("strategy", ["halo", "recompute"])
def test_decomposed_equals_whole_227(world, strategy):
    cfg = default_config(init=INIT_RAND, seed=11)
    x, w = make_problem(cfg)
    ref = forward_cpu(x, w, cfg)
    got = forward_cpu_decomposed(x, w, cfg, world, strategy)
    assert got.shape == ref.shape == (1, 13, 13, 256)
    assert torch.equal(got, ref)        # same op order per output => bit-identical


def test_rows_x_batch_plan_8_ranks():
    """8 ranks as 2 row groups x 4 batch groups: 7/6 final rows instead of 2,2,2,2,2,1,1,1; one halo neighbour."""
    plan = make_plan(default_config(batch=128), 8, "halo", row_groups=2)
    validate_plan(plan)
    assert (plan.row_groups, plan.batch_groups) == (2, 4)
    assert [rp.images for rp in plan.ranks] == [(0, 32)] * 2 + [(32, 64)] * 2 + [(64, 96)] * 2 + [(96, 128)] * 2
    assert [rp.out_rows for rp in plan.ranks] == [(0, 7), (7, 13)] * 4
    assert plan.ranks[5].halo == [(4, 12, 14)] and plan.ranks[4].halo == [(5, 14, 17)]
    assert plan.ranks[4].serves == [(5, 12, 14)]
    assert make_plan(default_config(batch=128), 8, "halo", "auto").row_groups == 2
    assert make_plan(default_config(batch=128), 3, "halo", "auto").row_groups == 3      # odd world: pure rows
    assert make_plan(default_config(batch=128), 8, "halo").row_groups == 8              # default: the reference's split
    with pytest.raises(ValueError):
        make_plan(default_config(batch=8), 8, "halo", row_groups=3)


@settings(max_examples=120, deadline=None)
@given(H=st.integers(35, 500), batch=st.integers(1, 9), R=st.integers(1, 4), B=st.integers(1, 4),
       strategy=st.sampled_from(["halo", "recompute"]))
def test_rows_x_batch_invariants(H, batch, R, B, strategy):
    validate_plan(make_plan(default_config(H=H, W=67, batch=batch), R * B, strategy, row_groups=R))


@pytest.mark.parametrize("world,R", [(2, 1), (4, 2), (8, 2), (6, 3), (8, 4)])
@pytest.mark.parametrize("strategy", ["halo", "recompute"])
def test_rows_x_batch_decomposed_equals_whole(world, R, strategy):
    cfg = default_config(H=131, W=99, batch=5, init=INIT_RAND, seed=3)     # 5 images over up to 4 batch groups: uneven
    x, w = make_problem(cfg)
    ref = forward_cpu(x, w, cfg)
    got = forward_cpu_decomposed(x, w, cfg, world, strategy, row_groups=R)
    assert torch.equal(got, ref)


def test_decomposed_more_ranks_than_rows():
    cfg = default_config(H=83, W=83, batch=2, init=INIT_RAND, seed=5)   # 4 final rows
    x, w = make_problem(cfg)
    ref = forward_cpu(x, w, cfg)
    assert ref.shape[1] == 4
    got = forward_cpu_decomposed(x, w, cfg, 6)
    assert torch.equal(got, ref)


def test_splitk_and_batch_shard_plans():
    """Host-side plans of the full-network extension (pure arithmetic, no GPU)."""
    from b200cnn.ops.ext import plan_splitk, effective_splitk
    # 81 k-blocks (K = 2592) asked in 10 splits run as 9 splits of 9: workspace and reduce must use 9
    assert effective_splitk(81, 10) == 9 and effective_splitk(81, 9) == 9 and effective_splitk(8, 16) == 8
    for nkb in range(1, 200):
        for ks in range(1, 40):
            e = effective_splitk(nkb, ks)
            assert effective_splitk(nkb, e) == e and e <= max(1, min(ks, nkb))
    assert plan_splitk(128, 256, 256, 81, 148) == effective_splitk(81, plan_splitk(128, 256, 256, 81, 148))
    from b200cnn.runners.full import batch_shard
    # FC6 at batch 256 on 148 SMs: 2 x 16 tiles of 128x256 -> 4 splits of 72 k-blocks
    assert plan_splitk(256, 4096, 256, 288, 148) == 4
    assert plan_splitk(256, 1000, 256, 128, 148) == 16          # FC8: 8 tiles, capped by 8 k-blocks per split
    assert plan_splitk(43264, 384, 256, 72, 148) == 1           # conv3 at batch 256 fills the GPU: no split
    assert plan_splitk(64, 128, 128, 8, 148) == 1               # too few k-blocks to split
    for batch, world in ((256, 8), (5, 2), (3, 8), (1, 4)):
        parts = [batch_shard(batch, r, world) for r in range(world)]
        assert parts[0][0] == 0 and parts[-1][1] == batch
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))     # contiguous, disjoint, complete
        assert all(0 <= b1 - b0 <= -(-batch // world) for b0, b1 in parts)


def _same_plan(cfg, world, strategy, R):
    from b200cnn.ops.cpu import native_plan_rank
    plan = make_plan(cfg, world, strategy, row_groups=R or None)
    for rp in plan.ranks:
        c = native_plan_rank(cfg, world, rp.rank, strategy, R)
        assert (c.row_group, c.batch_group, c.row_groups, c.batch_groups) == \
            (rp.row_group, rp.batch_group, plan.row_groups, plan.batch_groups)
        assert (c.img0, c.img1) == tuple(rp.images)
        for name, rng in (("out", rp.out_rows), ("conv2_", rp.conv2_rows), ("need", rp.pool1_need),
                          ("own", rp.pool1_own), ("conv1_", rp.conv1_rows), ("in", rp.in_rows)):
            got = (getattr(c, name + "0"), getattr(c, name + "1"))
            assert got == tuple(rng) or (got[1] <= got[0] and rng[1] <= rng[0]), (name, got, rng, rp.rank)
        assert [(h.peer, h.lo, h.hi) for h in c.halo[:c.n_halo]] == list(rp.halo)
        assert sorted((h.peer, h.lo, h.hi) for h in c.serves[:c.n_serves]) == sorted(rp.serves)
        sh = plan.shapes
        assert (c.Hc1, c.Hp1, c.Hc2, c.Hp2) == (sh.Hc1, sh.Hp1, sh.Hc2, sh.Hp2)


@pytest.mark.parametrize("world,R", [(1, 0), (2, 0), (3, 0), (4, 0), (8, 0), (4, 2), (8, 2), (8, 4), (6, 3), (16, 0)])
@pytest.mark.parametrize("strategy", ["halo", "recompute"])
def test_native_planner_equals_python_planner_227(world, R, strategy):
    """csrc/cpu/planner_cpu.cpp (the planner a native host uses) against parallel/planner.py, rank by rank."""
    _same_plan(default_config(batch=13), world, strategy, R)


@settings(max_examples=120, deadline=None)
@given(H=st.integers(min_value=40, max_value=700), batch=st.integers(min_value=0, max_value=9),
       R=st.integers(min_value=1, max_value=6), B=st.integers(min_value=1, max_value=3),
       strategy=st.sampled_from(["halo", "recompute"]))
def test_native_planner_equals_python_planner_any_shape(H, batch, R, B, strategy):
    _same_plan(default_config(H=H, W=max(H, 67), batch=batch), R * B, strategy, R)
