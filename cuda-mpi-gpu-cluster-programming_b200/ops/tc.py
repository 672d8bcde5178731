"""tcgen05 tensor-core conv blocks (csrc/cuda/tc_block.cu): TF32 operands, fp32 accumulate in TMEM.

``TcConv`` owns the device copy of one layer's weights in the K-major, TF32-rounded, k-padded
layout the kernel's TMA descriptor points at (created once; the reference re-uploads weights on
every call, ref v4_mpi_cuda/src/alexnet_mpi_cuda.cu:188-191)."""
from __future__ import annotations

import ctypes as C

import torch

from . import cuda as cu
from ._ctypes import fptr

_bound = False


def _lib():
    global _bound
    L = cu.lib()
    if not _bound:
        if not hasattr(L, "b200_tc_block"):
            raise ImportError("libb200cnn_cuda.so was built without tc_block.cu")
        L.b200_tc_conv_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.b200_tc_conv_destroy.argtypes = [C.c_void_p]
        L.b200_tc_block_tiles.argtypes = [C.POINTER(cu.BlockGeom)]
        L.b200_tc_block.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(cu.BlockGeom),
                                    C.c_int, C.c_void_p]
        L.b200_tc_block_tiles_rows.argtypes = [C.POINTER(cu.BlockGeom), C.c_int, C.c_int]
        L.b200_tc_set_policy.argtypes = [C.c_char_p, C.c_char_p]
        _bound = True
    return L


def set_policy(**kw) -> None:
    """Process-wide kernel choice (csrc/cuda/tc_block.cu ``Policy``): ``conv2="auto"|"single"|"db"|"pair"``,
    ``conv1="auto"|"cpasync"``, ``no_swap`` / ``no_direct`` / ``dbg`` integers, ``pdl=0|1`` (programmatic
    dependent launch of block 2 behind block 1).  Every choice gives bit-identical
    outputs; tests and experiments use it to compare the kernels."""
    for k, v in kw.items():
        cu.cu_check(_lib().b200_tc_set_policy(k.encode(), str(v).encode()), f"tc_set_policy({k}={v})")


class TcConv:
    def __init__(self, w: torch.Tensor, device):
        w = w.detach().cpu().contiguous().float()
        self.K, self.C, self.F, _ = w.shape
        self.device = torch.device(device)
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            cu.cu_check(_lib().b200_tc_conv_create(fptr(w), self.K, self.C, self.F, C.byref(h)),
                        "tc_conv_create")
        self._h = h

    def run(self, bias: torch.Tensor, out_ptr: int, g: cu.BlockGeom, round_out: bool = False,
            overlap_prev: bool = False) -> None:
        """``overlap_prev``: programmatic dependent launch — the kernel's prologue may overlap the tail of the
        previous kernel in the stream; it waits for that kernel before reading what it wrote (flags bit 1)."""
        cu.cu_check(_lib().b200_tc_block(self._h, fptr(bias), C.c_void_p(out_ptr), C.byref(g),
                                         (1 if round_out else 0) | (2 if overlap_prev else 0),
                                         cu.current_stream_ptr()), "tc_block")

    def __del__(self):
        try:
            if self._h:
                _lib().b200_tc_conv_destroy(self._h)
                self._h = None
        except Exception:
            pass


def block_tiles(g: cu.BlockGeom) -> int:
    n = int(_lib().b200_tc_block_tiles(C.byref(g)))
    if n < 0:
        cu.cu_check(n, "tc_block_tiles")
    return n


def block_tiles_rows(g: cu.BlockGeom, lo: int, hi: int) -> int:
    """Tiles per image that produce a pooled row of [lo, hi): arrival target of a reader of those rows."""
    n = int(_lib().b200_tc_block_tiles_rows(C.byref(g), lo, hi))
    if n < 0:
        cu.cu_check(n, "tc_block_tiles_rows")
    return n
