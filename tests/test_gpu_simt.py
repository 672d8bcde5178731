"""Exact-FP32 SIMT fused kernels on a real GPU vs the fp32 oracle; multi-rank emulation."""
import pytest
import torch

from b200cnn import default_config
from b200cnn.config import INIT_RAND, LRN_V1, LRN_V3
from b200cnn.models.init import make_problem
from b200cnn.models.alexnet import forward_torch_reference

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a - b).abs().max() / b.abs().max()).item()


def test_simt_const_closed_form():
    from b200cnn.models.engine import single_gpu_engine
    for mode, want in ((LRN_V1, 44.4152), (LRN_V3, 29.2932)):
        cfg = default_config(lrn_mode=mode, precision="simt")
        x, w = make_problem(cfg)
        eng = single_gpu_engine(cfg, w, "cuda:0")
        y = eng.forward(x.cuda()).cpu()
        eng.check_status()
        assert tuple(y.shape) == (1, 13, 13, 256)
        assert abs(y[0, 0, 0, 0].item() - want) < 2e-3
        assert abs(y[0, 12, 12, 0].item() - want) < 2e-3


@pytest.mark.parametrize("H,W,batch", [(227, 227, 3), (99, 131, 2), (67, 67, 1)])
def test_simt_matches_oracle(H, W, batch):
    from b200cnn.models.engine import single_gpu_engine
    cfg = default_config(H=H, W=W, batch=batch, init=INIT_RAND, seed=5, precision="simt")
    x, w = make_problem(cfg)
    ref = forward_torch_reference(x, w, cfg)
    eng = single_gpu_engine(cfg, w, "cuda:0")
    y = eng.forward(x.cuda()).cpu()
    eng.check_status()
    assert y.shape == ref.shape
    assert _rel(y, ref) < 1e-4
    # pool1 intermediate too
    p1 = eng.pool1_view().cpu()
    assert _rel(p1, forward_torch_reference(x, w, cfg, upto="pool1")) < 1e-4
    # repeat forward: resident buffers, same answer
    y2 = eng.forward(x.cuda()).cpu()
    assert torch.equal(y, y2)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("strategy,scatter", [("halo", "p2p"), ("halo", "sharded"), ("recompute", "p2p"),
                                              ("recompute", "sharded")])
def test_simt_emulated_ranks_equal_single(world, strategy, scatter):
    from b200cnn.models.engine import single_gpu_engine, EmulatedWorld
    cfg = default_config(batch=2, init=INIT_RAND, seed=9, precision="simt")
    x, w = make_problem(cfg)
    one = single_gpu_engine(cfg, w, "cuda:0").forward(x.cuda()).clone()
    ew = EmulatedWorld(cfg, w, world, "cuda:0", strategy=strategy, scatter=scatter, halo="pull")
    y = ew.forward(x.cuda())
    assert torch.equal(y, one)
    y = ew.forward(x.cuda())          # second epoch: counters keep growing, targets follow
    assert torch.equal(y, one)


def test_flag_kernels_and_watchdog():
    from b200cnn.ops import cuda as cu
    flags = torch.zeros(8, dtype=torch.int32, device="cuda")
    p = flags.data_ptr()
    cu.flag_signal([p, p + 4, p + 4], add=3)
    cu.flag_wait([p, p + 4], [3, 6], p + 28, 1_000_000_000)
    torch.cuda.synchronize()
    assert flags[:2].tolist() == [3, 6] and flags[7].item() == 0
    # unreachable target: watchdog must fire (20 ms) and latch TIMEOUT instead of hanging
    cu.flag_wait([p + 8], [1], p + 28, 20_000_000)
    torch.cuda.synchronize()
    assert flags[7].item() == -4


@pytest.mark.parametrize("world,R", [(2, 1), (4, 2), (8, 2), (8, 4), (6, 3)])
@pytest.mark.parametrize("halo", ["push", "pull"])
@pytest.mark.parametrize("scatter,gather", [("p2p", "rank0"), ("sharded", "rank0"), ("sharded", "local")])
def test_simt_rows_x_batch_equal_single(world, R, scatter, gather, halo):
    """R row groups x B batch groups (uneven: 5 images over up to 4 groups), flow-control gate instead of a
    barrier, pool1 slabs alternating by forward parity: three forwards in a row, bit-equal to one rank."""
    from b200cnn.models.engine import single_gpu_engine, EmulatedWorld
    cfg = default_config(batch=5, init=INIT_RAND, seed=13, precision="simt")
    x, w = make_problem(cfg)
    one = single_gpu_engine(cfg, w, "cuda:0").forward(x.cuda()).clone()
    ew = EmulatedWorld(cfg, w, world, "cuda:0", strategy="halo", scatter=scatter, row_groups=R, gather=gather,
                       halo=halo)
    for _ in range(3):
        assert torch.equal(ew.forward(x.cuda()), one)
