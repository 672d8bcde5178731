"""ctypes mirrors of the structs in include/b200cnn.h (single source of truth for layout)."""
from __future__ import annotations

import ctypes as C


class WinGeom(C.Structure):
    _fields_ = [(n, C.c_int) for n in (
        "N", "H", "W", "C", "in_base", "in_slab", "K", "F", "S", "P",
        "Ho", "Wo", "out_base", "out_slab", "o0", "o1")]


class LrnParams(C.Structure):
    _fields_ = [("N_lrn", C.c_int), ("alpha", C.c_float), ("beta", C.c_float),
                ("k", C.c_float), ("mode", C.c_int)]


class BlockGeom(C.Structure):
    _fields_ = [("conv", WinGeom), ("Fp", C.c_int), ("Sp", C.c_int), ("Hp", C.c_int),
                ("Wp", C.c_int), ("pool_base", C.c_int), ("pool_slab", C.c_int),
                ("p0", C.c_int), ("p1", C.c_int), ("lrn", LrnParams)]


class GRand(C.Structure):
    _fields_ = [("r", C.c_int32 * 34), ("idx", C.c_int32)]


LRN_MODES = {"alpha_over_n": 0, "alpha_raw": 1}

STATUS = {0: "OK", -1: "bad argument", -2: "CUDA error", -3: "row outside slab",
          -4: "cross-GPU wait timed out", -5: "unsupported"}


class NativeError(RuntimeError):
    pass


def check(rc: int, what: str, detail: str = "") -> None:
    if rc != 0:
        raise NativeError(f"{what}: {STATUS.get(rc, rc)} {detail}".strip())


def fptr(t):
    """float* of a contiguous fp32 torch tensor (CPU or CUDA)."""
    import torch
    assert t.dtype == torch.float32 and t.is_contiguous(), (t.dtype, t.is_contiguous())
    return C.c_void_p(t.data_ptr())
