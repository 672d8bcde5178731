"""The four conv2-type kernels of csrc/cuda/tc_block.cu — single-buffered (one 256-pixel tile fills TMEM),
double-buffered (128-pixel tiles, own epilogue warps), the tcgen05 CTA pair that splits the pixels and the one that
splits the channels (cta_group::2, TMA-fed operands, LRN halo channels over DSMEM) —
and the two conv1 loaders (pixel operand by TMA / by cp.async) accumulate every output in the same K order, so all
choices must be bit-identical.  A mismatch fails."""
import pytest
import torch

from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem

pytestmark = pytest.mark.gpu


def _forward(cfg, x, w, **policy):
    from b200cnn.models.engine import single_gpu_engine
    from b200cnn.ops import tc
    tc.set_policy(**policy)
    try:
        e = single_gpu_engine(cfg, w, "cuda:0", "tf32")
        e.use_graph = False
        y = e.forward(x.cuda()).clone()
        y2 = e.forward(x.cuda()).clone()          # second forward: barrier phases and TMEM buffers wrap
        e.check_status()
    finally:
        tc.set_policy(conv2="auto", conv1="auto", dbg=0)
    return y.cpu(), y2.cpu()


@pytest.mark.parametrize("batch,hw", [(6, 227), (40, 227), (3, 99), (5, 163)])
def test_conv2_kernels_bit_identical(batch, hw):
    cfg = default_config(H=hw, W=hw, batch=batch, init=INIT_RAND, seed=2, precision="tf32")
    x, w = make_problem(cfg)
    ref, ref2 = _forward(cfg, x, w, conv2="single")
    assert torch.equal(ref, ref2)
    for kind in ("db", "pair", "cs", "auto"):
        y, y2 = _forward(cfg, x, w, conv2=kind)
        assert torch.equal(y, ref), f"conv2={kind} differs: max |diff| {float((y - ref).abs().max()):.3e}"
        assert torch.equal(y2, ref), f"conv2={kind}: second forward differs"


@pytest.mark.parametrize("batch,hw", [(6, 227), (3, 99), (2, 131)])
def test_conv1_loaders_bit_identical(batch, hw):
    cfg = default_config(H=hw, W=hw, batch=batch, init=INIT_RAND, seed=4, precision="tf32")
    x, w = make_problem(cfg)
    a, a2 = _forward(cfg, x, w, conv1="cpasync")
    b, b2 = _forward(cfg, x, w, conv1="auto")
    assert torch.equal(a, a2) and torch.equal(b, b2)
    assert torch.equal(a, b), f"TMA-fed conv1 differs from cp.async-fed: {float((a - b).abs().max()):.3e}"


@pytest.mark.parametrize("batch,hw", [(6, 227), (3, 99), (2, 131), (5, 163)])
def test_conv1_all_tma_and_register_pooling_bit_identical(batch, hw):
    """conv1 with every k-block by TMA (5-D tail maps), a 4-stage ring and the 3x3 pooling done in registers
    against the same kernel with cp.async tails (dbg 1024) and / or the shared-memory strips (dbg 2048)."""
    cfg = default_config(H=hw, W=hw, batch=batch, init=INIT_RAND, seed=6, precision="tf32")
    x, w = make_problem(cfg)
    ref, ref2 = _forward(cfg, x, w, dbg=1024 | 2048)
    assert torch.equal(ref, ref2)
    for knob in (0, 1024, 2048):
        y, y2 = _forward(cfg, x, w, dbg=knob)
        assert torch.equal(y, ref), f"dbg={knob} differs: max |diff| {float((y - ref).abs().max()):.3e}"
        assert torch.equal(y2, ref), f"dbg={knob}: second forward differs"


@pytest.mark.parametrize("graph", [False, True])
def test_dependent_launch_bit_identical(graph):
    """Block 2 launched as a programmatic dependent of block 1 (prologue overlaps block 1's tail, then
    griddepcontrol.wait) against the plain stream-ordered launch, eager and as a replayed CUDA graph."""
    from b200cnn.models.engine import single_gpu_engine
    from b200cnn.ops import tc
    cfg = default_config(H=227, W=227, batch=24, init=INIT_RAND, seed=5, precision="tf32")
    x, w = make_problem(cfg)
    xd = x.cuda()
    outs = {}
    try:
        for pdl in (0, 1):
            tc.set_policy(pdl=pdl)
            e = single_gpu_engine(cfg, w, "cuda:0", "tf32")
            e.use_graph = graph
            ys = [e.forward(xd).clone() for _ in range(4)]       # graph: forwards 3 and 4 are replays
            e.check_status()
            assert all(torch.equal(ys[0], y) for y in ys[1:])
            outs[pdl] = ys[0].cpu()
    finally:
        tc.set_policy(pdl=1)
    assert torch.equal(outs[0], outs[1]), f"max |diff| {float((outs[0] - outs[1]).abs().max()):.3e}"
