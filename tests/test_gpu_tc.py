"""tcgen05 TF32 fused kernels vs the fp32 oracle (tolerances are TF32's: 10-bit mantissa operands,
fp32 accumulate; SURVEY §7.3 'exactness ladder')."""
import pytest
import torch

from b200cnn import default_config
from b200cnn.config import INIT_RAND, LRN_V1, LRN_V3
from b200cnn.models.init import make_problem
from b200cnn.models.alexnet import forward_torch_reference

pytestmark = pytest.mark.gpu
TOL = 4e-3          # max abs err / max abs ref


def _rel(a, b):
    return ((a - b).abs().max() / b.abs().max()).item()


def _rel_elementwise(a, b, floor=0.02):
    """Largest error relative to each element's own size (with a floor of 2 % of the largest output, below which TF32
    rounding of 2400 products is the noise level).  A global max-ratio alone would hide an error confined to
    small-magnitude outputs, e.g. one wrong border tap."""
    return ((a - b).abs() / (b.abs() + floor * b.abs().max())).max().item()


@pytest.mark.parametrize("H,W,batch", [(227, 227, 3), (99, 131, 2), (67, 67, 1), (163, 227, 5)])
def test_tc_matches_oracle(H, W, batch):
    from b200cnn.models.engine import single_gpu_engine
    cfg = default_config(H=H, W=W, batch=batch, init=INIT_RAND, seed=5, precision="tf32")
    x, w = make_problem(cfg)
    ref = forward_torch_reference(x, w, cfg)
    eng = single_gpu_engine(cfg, w, "cuda:0")
    y = eng.forward(x.cuda()).cpu()
    eng.check_status()
    p1 = eng.pool1_view().cpu()
    assert _rel(p1, forward_torch_reference(x, w, cfg, upto="pool1")) < TOL
    assert y.shape == ref.shape
    assert _rel(y, ref) < TOL
    assert _rel_elementwise(y, ref) < 5e-2
    # borders are where a wrong tap or a wrong pad would sit: check them on their own scale
    edge = torch.zeros_like(ref, dtype=torch.bool)
    edge[:, 0] = edge[:, -1] = True
    edge[:, :, 0] = edge[:, :, -1] = True
    assert ((y - ref).abs()[edge].max() / ref.abs()[edge].max()).item() < TOL
    y2 = eng.forward(x.cuda()).cpu()
    assert torch.equal(y, y2)


def test_tc_const_closed_form():
    from b200cnn.models.engine import single_gpu_engine
    for mode, want in ((LRN_V1, 44.4152), (LRN_V3, 29.2932)):
        cfg = default_config(lrn_mode=mode, precision="tf32")
        x, w = make_problem(cfg)
        eng = single_gpu_engine(cfg, w, "cuda:0")
        y = eng.forward(x.cuda()).cpu()
        assert abs(y[0, 0, 0, 0].item() - want) / want < 2e-3
        assert abs(y[0, 12, 12, 255].item() - want) / want < 2e-3


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("strategy,scatter", [("halo", "p2p"), ("halo", "sharded"), ("recompute", "p2p"),
                                              ("recompute", "sharded")])
def test_tc_emulated_ranks_equal_single(world, strategy, scatter):
    from b200cnn.models.engine import single_gpu_engine, EmulatedWorld
    cfg = default_config(batch=2, init=INIT_RAND, seed=9, precision="tf32")
    x, w = make_problem(cfg)
    one = single_gpu_engine(cfg, w, "cuda:0").forward(x.cuda()).clone()
    ew = EmulatedWorld(cfg, w, world, "cuda:0", strategy=strategy, scatter=scatter, halo="pull")
    for _ in range(2):
        y = ew.forward(x.cuda())
        # every output is the same K-ordered tensor-core accumulation wherever its tile lies
        assert torch.equal(y, one)


def test_tc_agrees_with_simt_twin():
    from b200cnn.models.engine import single_gpu_engine
    cfg = default_config(batch=2, init=INIT_RAND, seed=21)
    x, w = make_problem(cfg)
    a = single_gpu_engine(cfg, w, "cuda:0", "simt").forward(x.cuda()).clone()
    b = single_gpu_engine(cfg, w, "cuda:0", "tf32").forward(x.cuda()).clone()
    assert _rel(b, a) < TOL


@pytest.mark.parametrize("world,R", [(2, 1), (4, 2), (8, 2), (8, 4), (6, 3)])
@pytest.mark.parametrize("halo", ["push", "pull"])
@pytest.mark.parametrize("scatter,gather", [("p2p", "rank0"), ("sharded", "rank0"), ("sharded", "local")])
def test_tc_rows_x_batch_equal_single(world, R, scatter, gather, halo):
    """R row groups x B batch groups (uneven: 5 images over up to 4 groups), flow-control gate instead of a
    barrier, pool1 slabs alternating by forward parity: three forwards in a row, bit-equal to one rank."""
    from b200cnn.models.engine import single_gpu_engine, EmulatedWorld
    cfg = default_config(batch=5, init=INIT_RAND, seed=13, precision="tf32")
    x, w = make_problem(cfg)
    one = single_gpu_engine(cfg, w, "cuda:0").forward(x.cuda()).clone()
    ew = EmulatedWorld(cfg, w, world, "cuda:0", strategy="halo", scatter=scatter, row_groups=R, gather=gather,
                       halo=halo)
    for _ in range(3):
        assert torch.equal(ew.forward(x.cuda()), one)


@pytest.mark.parametrize("world", [2, 3, 8])
def test_tc_pure_rows_halo_push_equal_single(world):
    """Pure row split with pushed halos: at 8 ranks a one-row rank pushes to three neighbours."""
    from b200cnn.models.engine import single_gpu_engine, EmulatedWorld
    cfg = default_config(batch=3, init=INIT_RAND, seed=17, precision="tf32")
    x, w = make_problem(cfg)
    one = single_gpu_engine(cfg, w, "cuda:0").forward(x.cuda()).clone()
    ew = EmulatedWorld(cfg, w, world, "cuda:0", strategy="halo", scatter="sharded", halo="push")
    for _ in range(2):
        assert torch.equal(ew.forward(x.cuda()), one)
