"""CPU oracle: native layers vs a plain PyTorch fp32 restatement, closed-form constants
(SURVEY App. C), glibc rand() stream, both LRN conventions."""
import ctypes

import pytest
import torch

from b200cnn import default_config, block12_shapes
from b200cnn.config import LRN_V1, LRN_V3, INIT_RAND
from b200cnn.models.init import make_problem
from b200cnn.models.alexnet import forward_cpu, forward_torch_reference
from b200cnn.ops import cpu as cpu_ops


def test_shapes_227():
    sh = block12_shapes(default_config())
    assert (sh.Hc1, sh.Hp1, sh.Hc2, sh.Hp2) == (55, 27, 27, 13)
    assert sh.out_hwc == (13, 13, 256)
    assert sh.flops_per_image() == 2 * (3025 * 96 * 363 + 729 * 256 * 2400)


def test_shapes_2048():
    sh = block12_shapes(default_config(H=2048, W=2048))
    assert (sh.Hc1, sh.Hp1, sh.Hc2, sh.Hp2) == (510, 254, 254, 126)


def test_glibc_rand_stream_matches_libc():
    libc = ctypes.CDLL("libc.so.6")
    for seed in (1, 42, 1746812345):
        libc.srand(seed)
        want = [libc.rand() for _ in range(500)]
        g = cpu_ops.GlibcRand(seed)
        got = [g.next() for _ in range(500)]
        assert got == want


def test_const_init_closed_form_both_lrn():
    # App. C: V1/V2 convention 44.4152 42.4612 40.6967; V3/V4 29.2932 25.9153 23.3255
    for mode, want in ((LRN_V1, (44.4152, 42.4612, 40.6967)), (LRN_V3, (29.2932, 25.9153, 23.3255))):
        cfg = default_config(lrn_mode=mode)
        x, w = make_problem(cfg)
        y = forward_cpu(x, w, cfg)
        assert tuple(y.shape) == (1, 13, 13, 256)
        got = y[0, 0, 0, :3].tolist()
        for g_, w_ in zip(got, want):
            assert abs(g_ - w_) < 2e-3, (mode, got)
        # interior pixel equals the corner (pool2 == 87.12 everywhere)
        assert torch.allclose(y[0, 6, 6], y[0, 0, 0], rtol=1e-5)


@pytest.mark.parametrize("H,W,batch", [(67, 67, 2), (99, 83, 1)])
def test_native_matches_torch_small(H, W, batch):
    cfg = default_config(H=H, W=W, batch=batch, init=INIT_RAND, seed=7)
    x, w = make_problem(cfg)
    y = forward_cpu(x, w, cfg)
    ref = forward_torch_reference(x, w, cfg)
    assert y.shape == ref.shape
    assert torch.allclose(y, ref, rtol=1e-4, atol=1e-6), (y - ref).abs().max()


def test_native_matches_torch_227_rand():
    cfg = default_config(init=INIT_RAND, seed=3)
    x, w = make_problem(cfg)
    dims = []
    y = forward_cpu(x, w, cfg, trace=lambda *a: dims.append(a))
    ref = forward_torch_reference(x, w, cfg)
    assert torch.allclose(y, ref, rtol=1e-4, atol=1e-6)
    assert dims[0] == ("Input", 227, 227, 3) and dims[-1] == ("After LRN2", 13, 13, 256)
    # V1 random init gives outputs around 0.05-0.10
    assert 0.01 < float(y.mean()) < 0.2


def test_fnv1a64_known_vector():
    t = torch.tensor(list(b"hello"), dtype=torch.uint8)
    assert cpu_ops.fnv1a64(t) == 0xA430D84680AABD0B
