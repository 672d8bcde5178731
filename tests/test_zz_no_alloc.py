"""The allocation-free convolution entry point (b200_conv_nchw_noalloc), called through the C ABI.

Some users of the library cannot allocate device memory at call time — a server whose own pool has
reserved the whole GPU, or a latency-bound loop that cannot afford cudaMalloc's synchronisation —
so this entry point takes its scratch from an 8 MiB array built into the library.  The test checks
that it gives what the allocating entry point (b200_conv_nchw) gives, on two calls in a row that
reuse the shared scratch with different shapes, that it does not change the amount of free device
memory, and that an oversized problem is refused with a clear message.

The file name sorts last on purpose: if this test fails under `pytest -x`, every other test has
already run.
"""
import ctypes as C
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 1e-5        # same relative tolerance as the other fp32 convolution tests (tests/test_gpu_misc.py)


def _lib():
    from b200cnn.ops import _build
    L = C.CDLL(_build.build_cuda())                      # plain C ABI, no Python wrapper in between
    sig = [C.c_void_p] * 4 + [C.c_int] * 8 + [C.c_void_p]
    L.b200_conv_nchw.argtypes = sig
    L.b200_conv_nchw_noalloc.argtypes = sig
    L.b200_conv_nchw_noalloc_max_scratch_bytes.restype = C.c_longlong
    L.b200_conv_nchw_scratch_bytes.argtypes = [C.c_int] * 8
    L.b200_conv_nchw_scratch_bytes.restype = C.c_longlong
    L.b200_last_error.restype = C.c_char_p
    return L


def _call(fn, x, w, b, out, S, P):
    N, Cc, H, W = x.shape
    K, _, F, _ = w.shape
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    return fn(C.c_void_p(x.data_ptr()), C.c_void_p(w.data_ptr()), C.c_void_p(b.data_ptr()),
              C.c_void_p(out.data_ptr()), N, Cc, H, W, K, F, S, P, stream)


def _case(shape, seed):
    N, Cc, H, W, K, F, S, P = shape
    g = torch.Generator().manual_seed(seed)
    x = torch.randn((N, Cc, H, W), generator=g).cuda()
    w = (torch.randn((K, Cc, F, F), generator=g) * 0.05).cuda()
    b = torch.randn(K, generator=g).cuda()
    Ho, Wo = (H - F + 2 * P) // S + 1, (W - F + 2 * P) // S + 1
    return x, w, b, (N, K, Ho, Wo), S, P


def test_noalloc_matches_allocating_entry_point_twice_in_a_row():
    L = _lib()
    assert L.b200_conv_nchw_noalloc_max_scratch_bytes() == 8 << 20
    # two different problems back to back: the second one starts with the first one's leftovers in
    # the shared scratch (conv2-shaped, then conv1-shaped: different sub-buffer boundaries)
    cases = [_case((1, 96, 27, 27, 256, 5, 1, 2), 11), _case((1, 3, 227, 227, 96, 11, 4, 0), 12)]
    free_before = None
    for i, (x, w, b, oshape, S, P) in enumerate(cases):
        want = torch.empty(oshape, device="cuda")
        got = torch.full(oshape, float("nan"), device="cuda")
        assert _call(L.b200_conv_nchw, x, w, b, want, S, P) == 0, L.b200_last_error()
        torch.cuda.synchronize()
        if i == 1:
            # by now the library's device image (the 8 MiB array included) is set up on this GPU: case 0 was the
            # warm-up call the header asks for, so from here on free memory must not move
            free_before = torch.cuda.mem_get_info()[0]
        assert _call(L.b200_conv_nchw_noalloc, x, w, b, got, S, P) == 0, L.b200_last_error()
        torch.cuda.synchronize()
        if i == 1:
            assert torch.cuda.mem_get_info()[0] == free_before      # nothing was allocated or freed
        err = ((got - want).abs().max() / want.abs().max()).item()
        assert err < TOL, (i, err)
        # and against an independent fp32 reference (on the CPU: cuDNN would silently use TF32)
        ref = torch.nn.functional.conv2d(x.cpu(), w.cpu(), b.cpu(), stride=S, padding=P)
        assert ((got.cpu() - ref).abs().max() / ref.abs().max()).item() < TOL


def test_noalloc_rejects_problems_beyond_the_built_in_scratch():
    L = _lib()
    shape = (16, 96, 27, 27, 256, 5, 1, 2)                  # batch 16 of conv2: ~19 MB of scratch
    assert L.b200_conv_nchw_scratch_bytes(*shape) > L.b200_conv_nchw_noalloc_max_scratch_bytes()
    x, w, b, oshape, S, P = _case(shape, 13)
    out = torch.zeros(oshape, device="cuda")
    rc = _call(L.b200_conv_nchw_noalloc, x, w, b, out, S, P)
    assert rc == -3                                           # B200_ERR_RANGE
    msg = L.b200_last_error().decode()
    assert "scratch" in msg and str(8 << 20) in msg
    assert float(out.abs().sum()) == 0.0                      # refused before touching the output
