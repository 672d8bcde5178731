"""Full AlexNet (extension; BASELINE.json config 5): blocks 1-2, Conv3-5, Pool5, FC6-8, softmax.

The reference implements blocks 1-2 only; the rest exists there as a plan
(ref final_project/PROBLEMS.txt:188-190, summary.md:37-43,68: Conv3 384@3x3 p1, Conv4 384, Conv5 256,
FC 9216 -> 4096 -> 4096 -> 1000).  Layer order keeps the reference's block structure (LRN after pool 2).

* blocks 1-2: the fused tcgen05 kernels of :mod:`b200cnn.models.engine` (spatially decomposable);
* Conv3-5: the tcgen05 GEMM (``csrc/cuda/tc_gemm.cu``) as an implicit GEMM: its loader gathers the im2col
  rows straight from the NHWC activation (B200_CONV_IMPLICIT=0 materialises them instead), bias+ReLU in
  its epilogue, Pool5;
* FC6-8: the same GEMM, tensor-parallel across ranks over peer memory (:class:`TPClassifier`):
    FC6 column-parallel, its loader pulls the features out of rank 0's buffer (all-gather fused);
    FC7 row-parallel, partial sums go straight into the owning rank's shard with red.add over NVLink
    (GEMM -> reduce-scatter fused); FC8 column-parallel, pulling every peer's FC7 shard while it
    multiplies and writing its logits columns straight into rank 0's buffer (gather fused).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Optional

import torch

from ..config import NetConfig
from ..ops import cuda as cu
from ..ops import ext
from ..parallel.symm import LocalSymm
from .engine import single_gpu_engine
from .init import Blocks12Weights


@dataclass
class FullWeights:
    blocks12: Blocks12Weights
    w3: torch.Tensor; b3: torch.Tensor        # [384,256,3,3]
    w4: torch.Tensor; b4: torch.Tensor        # [384,384,3,3]
    w5: torch.Tensor; b5: torch.Tensor        # [256,384,3,3]
    w6: torch.Tensor; b6: torch.Tensor        # [4096, 9216] over NHWC-flattened pool5 (6*6*256)
    w7: torch.Tensor; b7: torch.Tensor        # [4096, 4096]
    w8: torch.Tensor; b8: torch.Tensor        # [1000, 4096]


def make_full_weights(b12: Blocks12Weights, seed: int = 1, feat: int = 9216) -> FullWeights:
    g = torch.Generator().manual_seed(seed)

    def rnd(*shape, fan_in):
        return (torch.rand(shape, generator=g) - 0.5) * (2.0 / fan_in ** 0.5)
    return FullWeights(b12,
                       rnd(384, 256, 3, 3, fan_in=2304), torch.full((384,), 0.05),
                       rnd(384, 384, 3, 3, fan_in=3456), torch.full((384,), 0.05),
                       rnd(256, 384, 3, 3, fan_in=3456), torch.full((256,), 0.05),
                       rnd(4096, feat, fan_in=feat), torch.full((4096,), 0.05),
                       rnd(4096, 4096, fan_in=4096), torch.full((4096,), 0.05),
                       rnd(1000, 4096, fan_in=4096), torch.zeros(1000))


def forward_torch_reference_full(x: torch.Tensor, w: FullWeights, cfg: NetConfig) -> torch.Tensor:
    """fp32 PyTorch restatement (tests only): NHWC input -> [N,1000] probabilities."""
    import torch.nn.functional as Fn
    from .alexnet import forward_torch_reference
    y = forward_torch_reference(x, w.blocks12, cfg).permute(0, 3, 1, 2)
    y = Fn.relu(Fn.conv2d(y, w.w3, w.b3, padding=1))
    y = Fn.relu(Fn.conv2d(y, w.w4, w.b4, padding=1))
    y = Fn.relu(Fn.conv2d(y, w.w5, w.b5, padding=1))
    y = Fn.max_pool2d(y, 3, 2).permute(0, 2, 3, 1).reshape(x.shape[0], -1)      # NHWC flatten
    y = Fn.relu(Fn.linear(y, w.w6, w.b6))
    y = Fn.relu(Fn.linear(y, w.w7, w.b7))
    return torch.softmax(Fn.linear(y, w.w8, w.b8), dim=1)


class ConvGemm:
    """3x3 (or FxF) convolution on the tcgen05 GEMM with bias+ReLU in the epilogue.  ``implicit``: the
    GEMM's loader gathers the im2col rows straight from the NHWC activation (b200_tc_conv_gemm);
    otherwise the col matrix is materialised first (b200_im2col_nhwc), which costs a write and a read
    of F*F times the activation."""

    def __init__(self, w_kcff: torch.Tensor, bias: torch.Tensor, device, pad: int,
                 implicit: Optional[bool] = None):
        K, Cc, F, _ = w_kcff.shape
        self.K, self.C, self.F, self.P = K, Cc, F, pad
        self.lin = ext.TcLinear(w_kcff.permute(0, 2, 3, 1).reshape(K, F * F * Cc), device)
        self.bias = bias.to(device).contiguous()
        if implicit is None:
            implicit = os.environ.get("B200_CONV_IMPLICIT", "1") == "1"
        self.implicit = bool(implicit) and Cc % 32 == 0

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        if self.implicit:
            return self.lin.conv(x, self.F, 1, self.P, self.bias, relu=True, round_out=True)
        N, H, W, _ = x.shape
        col = ext.im2col(x, self.F, 1, self.P)
        Ho, Wo = H - self.F + 2 * self.P + 1, W - self.F + 2 * self.P + 1
        return self.lin(col, self.bias, relu=True, round_out=True).view(N, Ho, Wo, self.K)


class TPClassifier:
    """FC6-FC8 + softmax of one rank, tensor-parallel over ``world`` ranks (see module docstring)."""

    def __init__(self, w: FullWeights, symm, rank: int, world: int, batch: int, feat: int = 9216):
        self.rank, self.world, self.B, self.feat = rank, world, batch, feat
        self.symm = symm
        dev = self.device = torch.device(symm.device)
        H = 4096
        assert H % world == 0 and (H // world) % 32 == 0, "hidden size must split into 32-multiples"
        self.hs = hs = H // world
        self.out_splits = [t.numel() for t in torch.tensor_split(torch.arange(1000), world)]
        self.out_off = [sum(self.out_splits[:i]) for i in range(world)]
        r = rank
        self.fc6 = ext.TcLinear(w.w6[r * hs:(r + 1) * hs], dev)
        self.b6 = w.b6[r * hs:(r + 1) * hs].to(dev).contiguous()
        self.fc7 = [ext.TcLinear(w.w7[s * hs:(s + 1) * hs, r * hs:(r + 1) * hs].contiguous(), dev) for s in range(world)]
        self.b7 = w.b7[r * hs:(r + 1) * hs].to(dev).contiguous()
        o0, on = self.out_off[r], self.out_splits[r]
        self.fc8 = [ext.TcLinear(w.w8[o0:o0 + on, s * hs:(s + 1) * hs].contiguous(), dev) for s in range(world)]
        self.b8 = w.b8[o0:o0 + on].to(dev).contiguous()
        f4 = 4
        # features: every rank holds a full-size buffer and fills the rows of the images whose conv3-5 it ran
        # (``x5_sources``); by default everything sits on rank 0
        self.buf_x5 = symm.alloc("tp_x5", batch * feat * f4)
        self.x5_sources = [(0, 0, batch)]                  # (rank, first image, end image)
        self.buf_y6 = symm.alloc("tp_y6", batch * hs * f4)
        self.buf_y7 = symm.alloc("tp_y7", batch * hs * f4)
        # FC7 partial products of this rank for every output shard s: [world][batch][hs]; shard owners pull them
        self.buf_part = symm.alloc("tp_part", world * batch * hs * f4 if world > 1 else 256)
        self.buf_logits = symm.alloc("tp_logits", batch * 1000 * f4 if rank == 0 else 256)
        self.buf_bar = symm.alloc("tp_bar", (world + 4) * 4)
        self._bar_n = 0

    # -- device-side barrier on its own counters (one GPU per rank only) -------------------------
    def barrier(self) -> None:
        if self.world == 1:
            return
        self._bar_n += 1
        W = range(self.world)
        cu.flag_barrier([self.symm.ptr(r, "tp_bar") + 4 * self.rank for r in W],
                        [self.symm.ptr(self.rank, "tp_bar") + 4 * r for r in W],
                        [self._bar_n] * self.world, self.symm.ptr(self.rank, "tp_bar") + 4 * self.world)

    def x5_view(self) -> torch.Tensor:
        return self.buf_x5.tensor((self.B, self.feat))

    # -- phases (a real rank runs them back to back with barriers; the emulation interleaves ranks) --
    def phase_fc6(self) -> None:
        """All-gather fused into the GEMM's loader: the feature rows are read straight out of the buffers of the
        ranks that produced them (one launch per source: the A operand is one pointer + row stride)."""
        for src, i0, i1 in self.x5_sources:
            if i1 > i0:
                self.fc6.run(self.symm.ptr(src, "tp_x5") + 4 * i0 * self.feat, self.feat, i1 - i0,
                             self.buf_y6.ptr + 4 * i0 * self.hs, self.hs, self.b6, ext.RELU | ext.ROUND_TF32)

    def phase_fc7(self) -> None:
        """Row-parallel FC7: this rank's slice of the hidden vector gives a partial product for EVERY output shard.
        They are stored locally (plain stores); the shard owners pull and add them in rank order
        (``phase_fc7_finish``) — a reduce-scatter without atomics, bit-reproducible."""
        if self.world == 1:              # nothing to reduce: bias + ReLU in the GEMM epilogue
            self.fc7[0].run(self.buf_y6.ptr, self.hs, self.B, self.buf_y7.ptr, self.hs, self.b7,
                            ext.RELU | ext.ROUND_TF32)
            return
        for s in range(self.world):
            self.fc7[s].run(self.buf_y6.ptr, self.hs, self.B, self.buf_part.ptr + 4 * s * self.B * self.hs, self.hs,
                            None, 0)

    def phase_fc7_finish(self) -> None:
        if self.world == 1:
            return
        mine = 4 * self.rank * self.B * self.hs
        ext.peer_reduce_([self.symm.ptr(k, "tp_part") + mine for k in range(self.world)], self.B, self.hs, self.hs,
                         self.b7, self.buf_y7.ptr, self.hs, ext.RELU | ext.ROUND_TF32)

    def phase_fc8(self) -> None:         # pulls every peer's shard; writes its logits columns on rank 0
        c = self.symm.ptr(0, "tp_logits") + 4 * self.out_off[self.rank]
        for s in range(self.world):
            self.fc8[s].run(self.symm.ptr(s, "tp_y7"), self.hs, self.B, c, 1000,
                            self.b8 if s == 0 else None, 0 if s == 0 else ext.ACCUMULATE)

    def phase_softmax(self) -> Optional[torch.Tensor]:
        if self.rank != 0:
            return None
        return ext.softmax(self.buf_logits.tensor((self.B, 1000)))

    def forward(self) -> Optional[torch.Tensor]:
        """Features must already be in rank 0's ``x5_view()``."""
        with torch.cuda.device(self.device):
            self.barrier()               # every rank's feature rows have landed
            self.phase_fc6()
            self.phase_fc7(); self.barrier()
            self.phase_fc7_finish(); self.barrier()
            self.phase_fc8(); self.barrier()
            return self.phase_softmax()


class FullAlexNet:
    """Single-process model: blocks 1-2 engine + conv3-5 + a ``world``-way tensor-parallel classifier
    whose ranks all live on this device (world=1: plain; world>1: single-GPU emulation of the TP
    protocol, launched phase by phase)."""

    def __init__(self, cfg: NetConfig, w: FullWeights, device="cuda:0", tp_world: int = 1):
        self.cfg = cfg
        self.device = torch.device(device)
        self.engine = single_gpu_engine(cfg, w.blocks12, device, cfg.precision)
        self.c3 = ConvGemm(w.w3, w.b3, device, 1)
        self.c4 = ConvGemm(w.w4, w.b4, device, 1)
        self.c5 = ConvGemm(w.w5, w.b5, device, 1)
        sh = self.engine.sh
        hp, wp = (sh.Hp2 - 3) // 2 + 1, (sh.Wp2 - 3) // 2 + 1
        self.feat = hp * wp * 256
        assert w.w6.shape[1] == self.feat, (w.w6.shape, self.feat)
        self.symm = LocalSymm(tp_world, device)
        self.tp: List[TPClassifier] = [TPClassifier(w, self.symm.view(r), r, tp_world, cfg.batch, self.feat)
                                       for r in range(tp_world)]

    def features(self, x: torch.Tensor) -> torch.Tensor:
        y = self.engine.forward(x)
        y = self.c5(self.c4(self.c3(y)))
        return ext.maxpool(y, 3, 2).reshape(x.shape[0], -1)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        with torch.cuda.device(self.device):
            self.tp[0].x5_view().copy_(self.features(x))
            T = self.tp
            for t in T: t.phase_fc6()
            for t in T: t.phase_fc7()
            for t in T: t.phase_fc7_finish()
            for t in T: t.phase_fc8()
            return T[0].phase_softmax()
