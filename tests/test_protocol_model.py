"""The barrier-free forward protocol, checked exhaustively on a model (parallel/protocol_model.py): every
interleaving of 2-4 ranks over 4-5 forwards is hazard- and deadlock-free for both halo modes, and taking away any
one ingredient (go gate, parity double buffer, arrival wait, rank 0's completion wait) is caught."""
import pytest

from b200cnn.parallel.protocol_model import check


@pytest.mark.parametrize("halo", ["push", "pull"])
@pytest.mark.parametrize("world,forwards", [(2, 5), (3, 4), (4, 4)])
def test_protocol_is_safe(world, forwards, halo):
    assert check(world, forwards, halo) is None


@pytest.mark.parametrize("halo", ["push", "pull"])
@pytest.mark.parametrize("mutation", ["no_go_gate", "single_slab", "no_arrival_wait", "no_done_wait"])
def test_every_ingredient_is_needed(mutation, halo):
    f = check(3, 4, halo, frozenset([mutation]))
    assert f is not None and f.kind == "hazard", (mutation, halo, f)
    assert f.trace and "region" in f.detail


def test_one_rank_has_nothing_to_wait_for():
    assert check(1, 4, "push") is None
