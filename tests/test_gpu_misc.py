"""GPU tests of the auxiliary native entry points: NCHW conv adaptor (conv_test) and HW1 DGEMM."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(1, 3, 32, 32, 8, 3, 1, 0), (2, 3, 227, 227, 96, 11, 4, 0),
                                   (2, 96, 27, 27, 256, 5, 1, 2), (3, 5, 19, 23, 7, 3, 2, 1)])
def test_conv_nchw_matches_torch(shape):
    from b200cnn.ops import cuda as cu
    N, C, H, W, K, F, S, P = shape
    g = torch.Generator().manual_seed(1)
    x = torch.randn((N, C, H, W), generator=g)
    w = torch.randn((K, C, F, F), generator=g) * 0.05
    b = torch.randn(K, generator=g)
    ref = torch.nn.functional.conv2d(x, w, b, stride=S, padding=P)
    out = cu.conv_nchw(x.cuda(), w.cuda(), b.cuda(), S, P).cpu()
    assert out.shape == ref.shape
    assert ((out - ref).abs().max() / ref.abs().max()).item() < 1e-5
    out2 = cu.conv_nchw(x.cuda(), w.cuda(), None, S, P).cpu()          # bias optional
    ref2 = torch.nn.functional.conv2d(x, w, None, stride=S, padding=P)
    assert ((out2 - ref2).abs().max() / ref2.abs().max()).item() < 1e-5


@pytest.mark.parametrize("shape", [(1, 3, 32, 32, 8, 3, 1, 0), (2, 3, 227, 227, 96, 11, 4, 0),
                                   (2, 96, 27, 27, 256, 5, 1, 2), (3, 5, 19, 23, 7, 3, 2, 1)])
def test_conv_test_tensor_core_path_matches_torch(shape):
    """conv_test's default GPU path: NCHW adaptor around the tcgen05 implicit-GEMM convolution."""
    from b200cnn.runners.conv_test import conv_nchw_tc
    N, C, H, W, K, F, S, P = shape
    g = torch.Generator().manual_seed(1)
    x = torch.randn((N, C, H, W), generator=g)
    w = torch.randn((K, C, F, F), generator=g) * 0.05
    b = torch.randn(K, generator=g)
    ref = torch.nn.functional.conv2d(x, w, b, stride=S, padding=P)
    out = conv_nchw_tc(x.cuda(), w, b, S, P).cpu()
    assert out.shape == ref.shape
    assert ((out - ref).abs().max() / ref.abs().max()).item() < 4e-3          # TF32 operands, fp32 accumulate


def test_conv_test_runner_prints_the_reference_line():
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for extra in ([], ["--exact"]):
        p = subprocess.run([sys.executable, "-m", "b200cnn.runners.conv_test", *extra], capture_output=True, text=True,
                           cwd=root, timeout=300)
        assert p.returncode == 0, p.stdout + p.stderr
        assert "Convolution Test Output (first 10 values):" in p.stdout and "Convolution Test: PASSED" in p.stdout


@pytest.mark.parametrize("rows,n", [(64, 64), (128, 256), (96, 192), (512, 1024)])
def test_dgemm_rows_exact_on_small_integers(rows, n):
    from b200cnn.ops import cuda as cu
    g = torch.Generator().manual_seed(2)
    A = torch.randint(0, 10, (rows, n), generator=g).double()
    B = torch.randint(0, 10, (n, n), generator=g).double()
    dA, dB = A.cuda(), B.cuda()
    dC = torch.empty((rows, n), dtype=torch.float64, device="cuda")
    cu.dgemm_rows(dA.data_ptr(), dB.data_ptr(), dC.data_ptr(), rows, n)
    torch.cuda.synchronize()
    assert torch.equal(dC.cpu(), A @ B)          # integer-valued inputs: FP64 result is exact


def test_hw1_runner_gpu_backend():
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, "-m", "b200cnn.runners.hw1", "512", "--backend", "gpu"],
                       capture_output=True, text=True, cwd=root, timeout=300)
    assert p.returncode == 0 and "Test: PASSED" in p.stdout, p.stdout + p.stderr
