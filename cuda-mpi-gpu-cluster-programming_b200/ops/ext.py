"""ctypes bindings for the full-AlexNet extension kernels (csrc/cuda/tc_gemm.cu, ext_ops.cu)."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from . import cuda as cu
from ._ctypes import fptr

_bound = False

RELU, ROUND_TF32, ACCUMULATE, BIAS = 1, 2, 4, 8


def _lib():
    global _bound
    L = cu.lib()
    if not _bound:
        L.b200_tc_gemm_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.b200_tc_gemm_destroy.argtypes = [C.c_void_p]
        L.b200_tc_gemm.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                   C.c_int, C.c_void_p]
        L.b200_tc_gemm_splitk.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                          C.c_int, C.c_int, C.c_longlong, C.c_void_p]
        L.b200_splitk_reduce.argtypes = [C.c_void_p, C.c_int, C.c_longlong, C.c_longlong, C.c_int, C.c_void_p,
                                         C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.b200_tc_gemm_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.b200_peer_reduce.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_longlong, C.c_int, C.c_int, C.c_void_p,
                                       C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.b200_tc_conv_gemm.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 7 + [C.c_void_p, C.c_void_p, C.c_int,
                                                                                 C.c_int, C.c_void_p]
        L.b200_im2col_nhwc.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 7 + [C.c_void_p]
        L.b200_maxpool_nhwc.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 6 + [C.c_void_p]
        L.b200_bias_act.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.b200_softmax_rows.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.b200_repitch_rows.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_void_p]
        L.b200_space_to_depth.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 9 + [C.c_void_p]
        _bound = True
    return L


def plan_splitk(M: int, N: int, bn: int, nkb: int, sms: int) -> int:
    """K splits for an [M, N] GEMM with 128 x ``bn`` tiles and ``nkb`` k-blocks of 32 on ``sms`` SMs: layers
    whose tiles fill less than half of the SMs are split along K until they do, each split keeping at
    least 8 k-blocks; layers with fewer than 16 k-blocks are never split."""
    tiles = ((M + 127) // 128) * ((N + bn - 1) // bn)
    if tiles * 2 > sms or nkb < 16:
        return 1
    return effective_splitk(nkb, max(1, min(sms // tiles, nkb // 8)))


def effective_splitk(nkb: int, ks: int) -> int:
    """The split count the kernel really runs for a request of ``ks``: every split covers ceil(nkb / ks) k-blocks,
    so fewer splits may be enough (81 k-blocks in 10 splits of 9 are 9 splits).  The workspace and the reduce must
    use THIS number — a reduce over the requested count would add a slab nobody wrote (and the C entry point
    refuses a count it would have to change)."""
    ks = max(1, min(int(ks), int(nkb)))
    kb_per = -(-nkb // ks)
    return -(-nkb // kb_per)


class TcLinear:
    """Device-resident weights [N,K] (nn.Linear layout) for b200_tc_gemm."""

    def __init__(self, w_nk: torch.Tensor, device):
        w = w_nk.detach().cpu().contiguous().float()
        self.N, self.K = w.shape
        self.device = torch.device(device)
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            cu.cu_check(_lib().b200_tc_gemm_create(fptr(w), self.N, self.K, C.byref(h)), "tc_gemm_create")
        self._h = h
        nkb, bn = C.c_int(), C.c_int()
        cu.cu_check(_lib().b200_tc_gemm_info(h, C.byref(nkb), C.byref(bn)), "tc_gemm_info")
        self.nkb, self.bn = nkb.value, bn.value
        self._sms = torch.cuda.get_device_properties(self.device).multi_processor_count
        self._ws: Optional[torch.Tensor] = None

    def splitk_for(self, M: int) -> int:
        """K splits for an M-row call: layers whose tiles fill less than half of the SMs (FC at batch <= 256)
        are split along K until they do; each split keeps >= 8 k-blocks.  B200_GEMM_SPLITK=0 disables."""
        if os.environ.get("B200_GEMM_SPLITK", "1") != "1":
            return 1
        return plan_splitk(M, self.N, self.bn, self.nkb, self._sms)

    def run(self, a_ptr: int, lda: int, M: int, c_ptr: int, ldc: int, bias: Optional[torch.Tensor] = None,
            flags: int = 0) -> None:
        if bias is not None:
            flags |= BIAS
        ks = 1 if (flags & ACCUMULATE) else effective_splitk(self.nkb, self.splitk_for(M))
        if ks > 1:                       # partial sums into a workspace, then a deterministic reduce + epilogue
            need = ks * M * self.N
            if self._ws is None or self._ws.numel() < need:
                self._ws = torch.empty(need, dtype=torch.float32, device=self.device)
            st = cu.current_stream_ptr()
            cu.cu_check(_lib().b200_tc_gemm_splitk(self._h, C.c_void_p(a_ptr), lda, M, None, fptr(self._ws), self.N,
                                                   0, ks, M * self.N, st), "tc_gemm_splitk")
            cu.cu_check(_lib().b200_splitk_reduce(fptr(self._ws), ks, M * self.N, M, self.N,
                                                  fptr(bias) if bias is not None else None, C.c_void_p(c_ptr), ldc,
                                                  flags, st), "splitk_reduce")
            return
        cu.cu_check(_lib().b200_tc_gemm(self._h, C.c_void_p(a_ptr), lda, M,
                                        fptr(bias) if bias is not None else None, C.c_void_p(c_ptr), ldc,
                                        flags, cu.current_stream_ptr()), "tc_gemm")

    def __call__(self, a: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False,
                 round_out: bool = False) -> torch.Tensor:
        assert a.dim() == 2 and a.shape[1] == self.K and a.is_contiguous()
        out = torch.empty((a.shape[0], self.N), dtype=torch.float32, device=a.device)
        self.run(a.data_ptr(), self.K, a.shape[0], out.data_ptr(), self.N, bias,
                 (RELU if relu else 0) | (ROUND_TF32 if round_out else 0))
        return out

    def conv(self, x: torch.Tensor, F: int, S: int, P: int, bias: Optional[torch.Tensor] = None,
             relu: bool = False, round_out: bool = False) -> torch.Tensor:
        """NHWC convolution as an implicit GEMM (weights [K_out, F*F*C] in (fy, fx, c) order): the loader
        gathers the im2col rows itself, nothing is materialised."""
        N, H, W, Cc = x.shape
        assert x.is_contiguous() and F * F * Cc == self.K and Cc % 32 == 0
        Ho, Wo = (H - F + 2 * P) // S + 1, (W - F + 2 * P) // S + 1
        out = torch.empty((N, Ho, Wo, self.N), dtype=torch.float32, device=x.device)
        flags = (RELU if relu else 0) | (ROUND_TF32 if round_out else 0) | (BIAS if bias is not None else 0)
        cu.cu_check(_lib().b200_tc_conv_gemm(self._h, fptr(x), N, H, W, Cc, F, S, P,
                                             fptr(bias) if bias is not None else None, fptr(out), self.N,
                                             flags, cu.current_stream_ptr()), "tc_conv_gemm")
        return out

    def __del__(self):
        try:
            if self._h:
                _lib().b200_tc_gemm_destroy(self._h)
                self._h = None
        except Exception:
            pass


def peer_reduce_(parts, rows: int, cols: int, ld_in: int, bias: Optional[torch.Tensor], out_ptr: int, ld: int,
                 flags: int = 0) -> None:
    """out[r, c] = act(sum_k parts[k][r, c] + bias[c]) with the partial matrices addressed by raw (possibly
    peer-mapped) pointers, added in list order: the deterministic reduce-scatter of the tensor-parallel FC7."""
    arr = (C.c_void_p * len(parts))(*parts)
    if bias is not None:
        flags |= BIAS
    cu.cu_check(_lib().b200_peer_reduce(arr, len(parts), rows, cols, ld_in, fptr(bias) if bias is not None else None,
                                        C.c_void_p(out_ptr), ld, flags, cu.current_stream_ptr()), "peer_reduce")


def im2col(x: torch.Tensor, F: int, S: int, P: int) -> torch.Tensor:
    N, H, W, Cc = x.shape
    Ho, Wo = (H - F + 2 * P) // S + 1, (W - F + 2 * P) // S + 1
    col = torch.empty((N * Ho * Wo, F * F * Cc), dtype=torch.float32, device=x.device)
    cu.cu_check(_lib().b200_im2col_nhwc(fptr(x), fptr(col), N, H, W, Cc, F, S, P, cu.current_stream_ptr()), "im2col")
    return col


def maxpool(x: torch.Tensor, F: int = 3, S: int = 2) -> torch.Tensor:
    N, H, W, Cc = x.shape
    out = torch.empty((N, (H - F) // S + 1, (W - F) // S + 1, Cc), dtype=torch.float32, device=x.device)
    cu.cu_check(_lib().b200_maxpool_nhwc(fptr(x), fptr(out), N, H, W, Cc, F, S, cu.current_stream_ptr()), "maxpool")
    return out


def bias_act_(x_ptr: int, bias: Optional[torch.Tensor], rows: int, cols: int, ld: int, relu: bool) -> None:
    cu.cu_check(_lib().b200_bias_act(C.c_void_p(x_ptr), fptr(bias) if bias is not None else None, rows, cols, ld,
                                     1 if relu else 0, cu.current_stream_ptr()), "bias_act")


def softmax(x: torch.Tensor) -> torch.Tensor:
    y = torch.empty_like(x)
    cu.cu_check(_lib().b200_softmax_rows(fptr(x), fptr(y), x.shape[0], x.shape[1], x.shape[1],
                                         cu.current_stream_ptr()), "softmax")
    return y


def space_to_depth_(in_ptr: int, out_ptr: int, N: int, H: int, W: int, Cc: int, S: int, in_base: int,
                    in_slab: int, y0: int, ys: int) -> None:
    cu.cu_check(_lib().b200_space_to_depth(C.c_void_p(in_ptr), C.c_void_p(out_ptr), N, H, W, Cc, S, in_base,
                                           in_slab, y0, ys, cu.current_stream_ptr()), "space_to_depth")


def s2d_weights(w: torch.Tensor, S: int) -> torch.Tensor:
    """[K,C,F,F] stride-S filter -> [K, S*S*C, Ft, Ft] stride-1 filter on the space-to-depth tensor
    (Ft = ceil(F/S); taps that fall outside the original FxF window are zero)."""
    K, Cc, F, _ = w.shape
    Ft = (F + S - 1) // S
    wp = torch.zeros((K, Cc, Ft * S, Ft * S), dtype=w.dtype)
    wp[:, :, :F, :F] = w
    # wp[k][c][ty*S+dy][tx*S+dx] -> out[k][(dy*S+dx)*C + c][ty][tx]
    wp = wp.view(K, Cc, Ft, S, Ft, S).permute(0, 3, 5, 1, 2, 4).reshape(K, S * S * Cc, Ft, Ft)
    return wp.contiguous()


def repitch_rows_(src_ptr: int, dst_ptr: int, rows: int, width: int, dpitch: int) -> None:
    cu.cu_check(_lib().b200_repitch_rows(C.c_void_p(src_ptr), C.c_void_p(dst_ptr), rows, width, dpitch,
                                         cu.current_stream_ptr()), "repitch_rows")
