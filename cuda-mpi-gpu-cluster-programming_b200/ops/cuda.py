"""ctypes bindings of libb200cnn_cuda.so (hand-written sm_100a kernels, include/b200cnn.h).

There is no PyTorch fallback: if the library cannot be loaded on a machine that has a GPU the
import of this module's ``lib()`` raises, so a silent eager-mode path can never masquerade as
the native one.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import torch

from . import _build
from ._ctypes import LrnParams, check, fptr, NativeError

MAX_SEG = 4
MAX_SIGNAL = 4
DEFAULT_TIMEOUT_NS = 5_000_000_000


class RowSeg(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("base", C.c_int), ("slab", C.c_int), ("lo", C.c_int),
                ("hi", C.c_int), ("flag", C.c_void_p), ("target", C.c_int), ("pad_", C.c_int)]


class RowMirror(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("base", C.c_int), ("slab", C.c_int), ("lo", C.c_int), ("hi", C.c_int)]


MAX_MIRROR = 4


class BlockGeom(C.Structure):
    _fields_ = [(n, C.c_int) for n in (
        "N", "H", "W", "C", "K", "F", "S", "P", "Hc", "Wc", "Hp", "Wp", "p0", "p1",
        "out_base", "out_slab")] + [
        ("lrn", LrnParams), ("nseg", C.c_int), ("n_signal", C.c_int), ("in_pitch", C.c_int),
        ("seg", RowSeg * MAX_SEG), ("signal", C.c_void_p * MAX_SIGNAL),
        ("status", C.c_void_p), ("timeout_ns", C.c_ulonglong), ("epoch", C.c_void_p),
        ("gate", C.c_void_p), ("gate_target", C.c_int), ("n_done", C.c_int),
        ("done_count", C.c_void_p), ("done", C.c_void_p * 2),
        ("signal_lo", C.c_int * MAX_SIGNAL), ("signal_hi", C.c_int * MAX_SIGNAL),
        ("n_mirror", C.c_int), ("pad2_", C.c_int), ("mirror", RowMirror * MAX_MIRROR)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_build.build_cuda())
        L.b200_last_error.restype = C.c_char_p
        L.b200_pack_weights_ffck.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.b200_simt_block.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.POINTER(BlockGeom), C.c_void_p]
        L.b200_simt_block_tiles.argtypes = [C.POINTER(BlockGeom)]
        L.b200_simt_block_tiles_rows.argtypes = [C.POINTER(BlockGeom), C.c_int, C.c_int]
        L.b200_symm_alloc.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]
        L.b200_symm_free.argtypes = [C.c_void_p]
        L.b200_ipc_export.argtypes = [C.c_void_p, C.c_char_p]
        L.b200_ipc_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
        L.b200_ipc_close.argtypes = [C.c_void_p]
        L.b200_can_access_peer.argtypes = [C.c_int, C.c_int]
        L.b200_enable_peer.argtypes = [C.c_int]
        L.b200_memcpy2d_h2d.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t,
                                        C.c_void_p]
        L.b200_memcpy2d_d2h.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t,
                                        C.c_void_p]
        L.b200_host_register.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.b200_host_unregister.argtypes = [C.c_void_p]
        L.b200_host_flag_store.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.b200_flag_signal.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_void_p]
        L.b200_flag_wait.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.c_int,
                                     C.c_void_p, C.c_ulonglong, C.c_void_p, C.c_void_p]
        L.b200_flag_begin.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_int, C.c_void_p]
        L.b200_flag_barrier.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_int),
                                        C.c_int, C.c_void_p, C.c_ulonglong, C.c_void_p, C.c_void_p]
        L.b200_conv_nchw.argtypes = [C.c_void_p] * 4 + [C.c_int] * 8 + [C.c_void_p]
        L.b200_conv_nchw_scratch_bytes.argtypes = [C.c_int] * 8
        L.b200_conv_nchw_scratch_bytes.restype = C.c_longlong
        L.b200_dgemm_rows.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        _lib = L
    return _lib


def cu_check(rc: int, what: str) -> None:
    if rc != 0:
        check(rc, what, (lib().b200_last_error() or b"").decode())


def current_stream_ptr() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def pack_weights_ffck(w: torch.Tensor) -> torch.Tensor:
    """[K,C,F,F] -> [F,F,C,K] (host), see csrc/cuda/common.cu."""
    w = w.detach().cpu().contiguous()
    K, Cc, F, _ = w.shape
    o = torch.empty((F, F, Cc, K), dtype=torch.float32)
    cu_check(lib().b200_pack_weights_ffck(fptr(w), fptr(o), K, Cc, F), "pack_weights_ffck")
    return o


def make_block_geom(*, N, H, W, Cin, K, F, S, P, Hc, Wc, Hp, Wp, rows: Tuple[int, int],
                    out_base: int, out_slab: int, lrn: Optional[LrnParams],
                    segs: Sequence[dict], signals: Sequence[int] = (),
                    status_ptr: int = 0, timeout_ns: int = DEFAULT_TIMEOUT_NS,
                    epoch_ptr: int = 0, in_pitch: int = 0, gate_ptr: int = 0, gate_target: int = 0,
                    done_count_ptr: int = 0, done: Sequence[int] = (),
                    mirrors: Sequence[dict] = (), signal_rows: Sequence[Tuple[int, int]] = ()) -> BlockGeom:
    g = BlockGeom()
    g.N, g.H, g.W, g.C = N, H, W, Cin
    g.K, g.F, g.S, g.P = K, F, S, P
    g.Hc, g.Wc, g.Hp, g.Wp = Hc, Wc, Hp, Wp
    g.p0, g.p1 = rows
    g.out_base, g.out_slab = out_base, out_slab
    g.lrn = lrn if lrn is not None else LrnParams(0, 0.0, 0.0, 0.0, 0)
    if not 1 <= len(segs) <= MAX_SEG:
        raise NativeError(f"need 1..{MAX_SEG} row segments, got {len(segs)}")
    if len(signals) > MAX_SIGNAL:
        raise NativeError(f"at most {MAX_SIGNAL} signal targets, got {len(signals)}")
    g.nseg = len(segs)
    for i, s in enumerate(segs):
        g.seg[i] = RowSeg(s["ptr"], s["base"], s["slab"], s["lo"], s["hi"],
                          s.get("flag", 0) or None, s.get("target", 0), 0)
    g.n_signal = len(signals)
    for i, p in enumerate(signals):
        g.signal[i] = p
        lo, hi = signal_rows[i] if i < len(signal_rows) else (0, 0)      # (0, 0): every tile signals
        g.signal_lo[i], g.signal_hi[i] = lo, hi
    g.status = status_ptr or None
    g.timeout_ns = timeout_ns
    g.epoch = epoch_ptr or None
    g.in_pitch = in_pitch
    g.gate = gate_ptr or None
    g.gate_target = gate_target
    if len(done) > 2:
        raise NativeError("at most 2 completion targets")
    g.done_count = (done_count_ptr or None) if done else None
    g.n_done = len(done)
    for i, p in enumerate(done):
        g.done[i] = p
    if len(mirrors) > MAX_MIRROR:
        raise NativeError(f"at most {MAX_MIRROR} halo-push targets, got {len(mirrors)}")
    g.n_mirror = len(mirrors)
    for i, m in enumerate(mirrors):
        g.mirror[i] = RowMirror(m["ptr"], m["base"], m["slab"], m["lo"], m["hi"])
    return g


def simt_block(w_ffck: torch.Tensor, bias: torch.Tensor, out_ptr: int, g: BlockGeom) -> None:
    cu_check(lib().b200_simt_block(fptr(w_ffck), fptr(bias), C.c_void_p(out_ptr), C.byref(g),
                                   current_stream_ptr()), "simt_block")


def simt_block_tiles(g: BlockGeom) -> int:
    return int(lib().b200_simt_block_tiles(C.byref(g)))


def simt_block_tiles_rows(g: BlockGeom, lo: int, hi: int) -> int:
    return int(lib().b200_simt_block_tiles_rows(C.byref(g), lo, hi))


def flag_signal(targets: Sequence[int], add: int = 1) -> None:
    if not targets:
        return
    arr = (C.c_void_p * len(targets))(*targets)
    cu_check(lib().b200_flag_signal(arr, len(targets), add, current_stream_ptr()), "flag_signal")


def flag_begin(epoch_ptr: int = 0, go_targets: Sequence[int] = ()) -> None:
    """Opens a forward on the current stream: ++*epoch (device) and +1 on every ``go`` counter."""
    n = len(go_targets)
    arr = (C.c_void_p * max(n, 1))(*go_targets)
    cu_check(lib().b200_flag_begin(C.c_void_p(epoch_ptr or None), arr, n, current_stream_ptr()), "flag_begin")


def flag_wait(flags: Sequence[int], targets: Sequence[int], status_ptr: int = 0,
              timeout_ns: int = DEFAULT_TIMEOUT_NS, epoch_ptr: int = 0) -> None:
    if not flags:
        return
    fa = (C.c_void_p * len(flags))(*flags)
    ta = (C.c_int * len(flags))(*targets)
    cu_check(lib().b200_flag_wait(fa, ta, len(flags), C.c_void_p(status_ptr or None),
                                  timeout_ns, C.c_void_p(epoch_ptr or None), current_stream_ptr()), "flag_wait")


def dgemm_rows(a_ptr: int, b_ptr: int, c_ptr: int, rows: int, n: int) -> None:
    """C[rows,n] = A[rows,n] @ B[n,n] in FP64 (pointers may be peer-mapped)."""
    cu_check(lib().b200_dgemm_rows(C.c_void_p(a_ptr), C.c_void_p(b_ptr), C.c_void_p(c_ptr), rows, n,
                                   current_stream_ptr()), "dgemm_rows")


def conv_nchw(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], S: int, P: int) -> torch.Tensor:
    """Plain convolution of an NCHW CUDA tensor with [K,C,F,F] weights -> NCHW (conv_test path)."""
    N, Cc, H, W = x.shape
    K, _, F, _ = w.shape
    Ho, Wo = (H - F + 2 * P) // S + 1, (W - F + 2 * P) // S + 1
    out = torch.empty((N, K, Ho, Wo), dtype=torch.float32, device=x.device)
    cu_check(lib().b200_conv_nchw(fptr(x.contiguous()), fptr(w.contiguous()),
                                  fptr(bias.contiguous()) if bias is not None else None, fptr(out),
                                  N, Cc, H, W, K, F, S, P, current_stream_ptr()), "conv_nchw")
    return out


def flag_barrier(signal_targets: Sequence[int], wait_flags: Sequence[int], wait_targets: Sequence[int],
                 status_ptr: int = 0, timeout_ns: int = DEFAULT_TIMEOUT_NS, epoch_ptr: int = 0) -> None:
    """Cross-GPU barrier in one launch (ranks must own one GPU each)."""
    n = len(signal_targets)
    sa = (C.c_void_p * n)(*signal_targets)
    fa = (C.c_void_p * n)(*wait_flags)
    ta = (C.c_int * n)(*wait_targets)
    cu_check(lib().b200_flag_barrier(sa, fa, ta, n, C.c_void_p(status_ptr or None), timeout_ns,
                                     C.c_void_p(epoch_ptr or None), current_stream_ptr()), "flag_barrier")


def memcpy2d_h2d(dst_ptr: int, dpitch: int, src: torch.Tensor, spitch: int, width: int, height: int) -> None:
    """Pitched pinned-host -> device copy on the current stream (bytes)."""
    cu_check(lib().b200_memcpy2d_h2d(C.c_void_p(dst_ptr), dpitch, C.c_void_p(src.data_ptr()), spitch, width,
                                     height, current_stream_ptr()), "memcpy2d_h2d")


def memcpy2d_d2h(dst_host_ptr: int, dpitch: int, src_ptr: int, spitch: int, width: int, height: int) -> None:
    """Pitched device -> pinned-host copy on the current stream (bytes)."""
    cu_check(lib().b200_memcpy2d_d2h(C.c_void_p(dst_host_ptr), dpitch, C.c_void_p(src_ptr), spitch, width,
                                     height, current_stream_ptr()), "memcpy2d_d2h")


def host_register(host_ptr: int, nbytes: int) -> int:
    """Pins foreign host memory (shared memory) and returns its device-side address."""
    d = C.c_void_p()
    cu_check(lib().b200_host_register(C.c_void_p(host_ptr), nbytes, C.byref(d)), "host_register")
    return int(d.value)


def host_unregister(host_ptr: int) -> None:
    lib().b200_host_unregister(C.c_void_p(host_ptr))


def host_flag_store(flag_dev_ptr: int, value: int) -> None:
    cu_check(lib().b200_host_flag_store(C.c_void_p(flag_dev_ptr), value, current_stream_ptr()), "host_flag_store")
