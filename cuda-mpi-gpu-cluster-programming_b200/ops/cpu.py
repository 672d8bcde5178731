"""CPU layer ops (native, csrc/cpu) on torch CPU tensors — the oracle and the V1/V2 compute.

Python-level signatures echo the reference's L2 layer API
(ref final_project/v1_serial/include/layers.hpp:9-42: serialConvLayer, serialReluLayer,
serialMaxPoolLayer, serialLRNLayer) with a batch dimension and row-slab arguments added.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import _build
from ._ctypes import WinGeom, LrnParams, GRand, LRN_MODES, check, fptr
from ..config import LayerParams
from ..shapes import conv_out_dim_or_zero, pool_out_dim_or_zero

_lib = None


class PlanBlock(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("F", "S", "P", "F_pool", "S_pool")]


class PlanHalo(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("peer", "lo", "hi")]


PLAN_MAX_HALO = 8


class RankPlanC(C.Structure):
    """``b200_rank_plan`` of include/b200cnn.h (the native planner's answer for one rank)."""
    _fields_ = [(n, C.c_int) for n in ("rank", "world", "row_group", "batch_group", "row_groups", "batch_groups",
                                       "img0", "img1", "out0", "out1", "conv2_0", "conv2_1", "need0", "need1",
                                       "own0", "own1", "conv1_0", "conv1_1", "in0", "in1", "n_halo", "n_serves")] + \
               [("halo", PlanHalo * PLAN_MAX_HALO), ("serves", PlanHalo * PLAN_MAX_HALO)] + \
               [(n, C.c_int) for n in ("Hc1", "Hp1", "Hc2", "Hp2")]


def native_plan_rank(cfg, world: int, rank: int, strategy: str = "halo", row_groups: int = 0) -> RankPlanC:
    """The native planner (csrc/cpu/planner_cpu.cpp) for one rank of ``world``; ``row_groups`` 0 = pure row split."""
    p1, p2 = cfg.conv1, cfg.conv2
    b1 = PlanBlock(p1.F, p1.S, p1.P, p1.F_pool, p1.S_pool)
    b2 = PlanBlock(p2.F, p2.S, p2.P, p2.F_pool, p2.S_pool)
    out = RankPlanC()
    rc = lib().b200_plan_rank(cfg.H, cfg.batch, C.byref(b1), C.byref(b2), world,
                              {"halo": 0, "recompute": 1}[strategy], int(row_groups or 0), rank, C.byref(out))
    if rc != 0:
        raise ValueError(f"b200_plan_rank failed with status {rc}")
    return out


def native_weights_read(path: str, name: str) -> torch.Tensor:
    """Tensor ``name`` of a weight file through the native reader (``b200_weights_read``): what a C/C++ host gets."""
    n, nd = C.c_size_t(0), C.c_int(0)
    shape = (C.c_int * 4)()
    rc = lib().b200_weights_read(path.encode(), name.encode(), None, 0, C.byref(n), shape, C.byref(nd))
    if rc != 0:
        raise ValueError(f"b200_weights_read({path}, {name}) failed with status {rc}")
    t = torch.empty(n.value, dtype=torch.float32)
    rc = lib().b200_weights_read(path.encode(), name.encode(), C.c_void_p(t.data_ptr()), n.value, C.byref(n), shape,
                                 C.byref(nd))
    if rc != 0:
        raise ValueError(f"b200_weights_read({path}, {name}) failed with status {rc}")
    return t.view(*[shape[i] for i in range(nd.value)]) if nd.value else t


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_build.build_cpu())
        L.b200_cpu_conv.argtypes = [C.c_void_p] * 4 + [C.POINTER(WinGeom)]
        L.b200_cpu_relu.argtypes = [C.c_void_p, C.c_size_t]
        L.b200_cpu_maxpool.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(WinGeom)]
        L.b200_cpu_lrn.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.POINTER(LrnParams)]
        L.b200_grand_seed.argtypes = [C.POINTER(GRand), C.c_uint32]
        L.b200_grand_seed.restype = None
        L.b200_grand_next.argtypes = [C.POINTER(GRand)]
        L.b200_grand_next.restype = C.c_int32
        for f in (L.b200_grand_fill_data, L.b200_grand_fill_weights):
            f.argtypes = [C.POINTER(GRand), C.c_void_p, C.c_size_t]
            f.restype = None
        L.b200_fnv1a64.argtypes = [C.c_void_p, C.c_size_t]
        L.b200_fnv1a64.restype = C.c_uint64
        L.b200_cpu_matmul_f64.argtypes = [C.c_void_p] * 3 + [C.c_int, C.c_int]
        L.b200_cpu_matmul_f64.restype = None
        L.b200_hw1_init_data.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.b200_hw1_init_data.restype = None
        L.b200_hw1_check.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_double]
        L.b200_weights_read.argtypes = [C.c_char_p, C.c_char_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t),
                                        C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.b200_plan_rank.argtypes = [C.c_int, C.c_int, C.POINTER(PlanBlock), C.POINTER(PlanBlock), C.c_int, C.c_int,
                                     C.c_int, C.c_int, C.POINTER(RankPlanC)]
        _lib = L
    return _lib


def _slab(t: torch.Tensor, base: Optional[int]):
    """(N, rows, W, C) slab tensor + global row of its first row."""
    assert t.dim() == 4, "expect NHWC"
    return t, (0 if base is None else base)


def conv(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, F: int, S: int, P: int, *,
         H: Optional[int] = None, in_base: int = 0,
         rows: Optional[Tuple[int, int]] = None) -> Tuple[torch.Tensor, int]:
    """Direct convolution of an NHWC slab.  ``H`` = logical full height (default: slab is
    the whole tensor).  ``rows`` = global output rows [o0,o1) to compute (default: all).
    Returns (out slab, global row of its first row)."""
    N, slab, W, Cin = x.shape
    H = slab if H is None else H
    K = w.shape[0]
    Ho, Wo = conv_out_dim_or_zero(H, F, P, S), conv_out_dim_or_zero(W, F, P, S)
    o0, o1 = (0, Ho) if rows is None else rows
    out = torch.empty((N, max(o1 - o0, 0), Wo, K), dtype=torch.float32)
    g = WinGeom(N, H, W, Cin, in_base, slab, K, F, S, P, Ho, Wo, o0, max(o1 - o0, 0), o0, o1)
    if out.numel():
        check(lib().b200_cpu_conv(fptr(x.contiguous()), fptr(w.contiguous()),
                                  fptr(b.contiguous()), fptr(out), C.byref(g)), "cpu conv")
    return out, o0


def relu_(x: torch.Tensor) -> torch.Tensor:
    check(lib().b200_cpu_relu(fptr(x), x.numel()), "cpu relu")
    return x


def maxpool(x: torch.Tensor, F: int, S: int, *, H: Optional[int] = None, in_base: int = 0,
            rows: Optional[Tuple[int, int]] = None) -> Tuple[torch.Tensor, int]:
    N, slab, W, Cc = x.shape
    H = slab if H is None else H
    Ho, Wo = pool_out_dim_or_zero(H, F, S), pool_out_dim_or_zero(W, F, S)
    o0, o1 = (0, Ho) if rows is None else rows
    out = torch.empty((N, max(o1 - o0, 0), Wo, Cc), dtype=torch.float32)
    g = WinGeom(N, H, W, Cc, in_base, slab, Cc, F, S, 0, Ho, Wo, o0, max(o1 - o0, 0), o0, o1)
    if out.numel():
        check(lib().b200_cpu_maxpool(fptr(x.contiguous()), fptr(out), C.byref(g)), "cpu maxpool")
    return out, o0


def lrn(x: torch.Tensor, p: LayerParams, mode: str = "alpha_over_n") -> torch.Tensor:
    out = torch.empty_like(x)
    lp = LrnParams(p.N_lrn, p.alpha, p.beta, p.k_lrn, LRN_MODES[mode])
    if x.numel():
        Cc = x.shape[-1]
        check(lib().b200_cpu_lrn(fptr(x.contiguous()), fptr(out), x.numel() // Cc, Cc,
                                 C.byref(lp)), "cpu lrn")
    return out


def fnv1a64(t: torch.Tensor) -> int:
    t = t.detach().cpu().contiguous()
    return int(lib().b200_fnv1a64(C.c_void_p(t.data_ptr()), t.numel() * t.element_size()))


class GlibcRand:
    """glibc ``srand``/``rand`` stream (see csrc/cpu/util_cpu.cpp)."""

    def __init__(self, seed: int):
        self.st = GRand()
        lib().b200_grand_seed(C.byref(self.st), seed & 0xFFFFFFFF)

    def next(self) -> int:
        return int(lib().b200_grand_next(C.byref(self.st)))

    def data(self, n: int) -> torch.Tensor:
        t = torch.empty(n, dtype=torch.float32)
        lib().b200_grand_fill_data(C.byref(self.st), fptr(t), n)
        return t

    def weights(self, n: int) -> torch.Tensor:
        t = torch.empty(n, dtype=torch.float32)
        lib().b200_grand_fill_weights(C.byref(self.st), fptr(t), n)
        return t
