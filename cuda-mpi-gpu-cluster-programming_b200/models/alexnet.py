"""AlexNet blocks 1-2 forward: Conv1->ReLU->Pool1->Conv2->ReLU->Pool2->LRN2.

Layer order is the reference's (LRN after pool, block 2 only —
ref v1_serial/src/alexnet_serial.cpp:89-163); do not "fix" it to canonical AlexNet.

This module holds the CPU paths (native oracle; exact row-decomposed pieces used by the V2.2
runner) and a plain-PyTorch fp32 restatement used only to cross-check the oracle in tests.
GPU paths live in :mod:`b200cnn.models.engine`.
"""
from __future__ import annotations

from typing import Callable, Optional

import torch

from ..config import NetConfig
from ..ops import cpu as cpu_ops
from ..parallel.planner import RankPlan
from .init import Blocks12Weights

TraceFn = Optional[Callable[[str, int, int, int], None]]


def forward_cpu(x: torch.Tensor, w: Blocks12Weights, cfg: NetConfig,
                trace: TraceFn = None) -> torch.Tensor:
    """Whole-image forward on CPU (V1 / V2.1 / V2.2-np1 semantics; ref
    v1_serial/src/alexnet_serial.cpp:67-186, v2_mpi_only/2.2_scatter_halo/src/alexnet_mpi.cpp:4-38).
    ``trace(layer, H, W, C)`` is called after every layer (V1 prints these)."""
    p1, p2 = cfg.conv1, cfg.conv2
    t = trace or (lambda *a: None)
    t("Input", x.shape[1], x.shape[2], x.shape[3])
    y, _ = cpu_ops.conv(x, w.w1, w.b1, p1.F, p1.S, p1.P)
    t("After Conv1", *y.shape[1:])
    cpu_ops.relu_(y)
    t("After ReLU1", *y.shape[1:])
    y, _ = cpu_ops.maxpool(y, p1.F_pool, p1.S_pool)
    t("After Pool1", *y.shape[1:])
    y, _ = cpu_ops.conv(y, w.w2, w.b2, p2.F, p2.S, p2.P)
    t("After Conv2", *y.shape[1:])
    cpu_ops.relu_(y)
    t("After ReLU2", *y.shape[1:])
    y, _ = cpu_ops.maxpool(y, p2.F_pool, p2.S_pool)
    t("After Pool2", *y.shape[1:])
    y = cpu_ops.lrn(y, p2, cfg.lrn_mode)
    t("After LRN2", *y.shape[1:])
    return y


def block1_rank_cpu(x_slab: torch.Tensor, in_base: int, w: Blocks12Weights, cfg: NetConfig,
                    rp: RankPlan, H: int) -> torch.Tensor:
    """Conv1->ReLU->Pool1 for the pool1 rows ``rp.pool1_own`` from the input slab
    ``[in_base, in_base+rows)``.  Returns the owned pool1 slab."""
    p1 = cfg.conv1
    y, c_base = cpu_ops.conv(x_slab, w.w1, w.b1, p1.F, p1.S, p1.P, H=H, in_base=in_base,
                             rows=rp.conv1_rows)
    cpu_ops.relu_(y)
    Hc1 = (H - p1.F + 2 * p1.P) // p1.S + 1
    y, _ = cpu_ops.maxpool(y, p1.F_pool, p1.S_pool, H=Hc1, in_base=c_base, rows=rp.pool1_own)
    return y


def block2_rank_cpu(pool1_slab: torch.Tensor, slab_base: int, w: Blocks12Weights,
                    cfg: NetConfig, rp: RankPlan, Hp1: int) -> torch.Tensor:
    """Conv2->ReLU->Pool2->LRN2 for final rows ``rp.out_rows`` from the pool1 slab that holds
    ``rp.pool1_need`` (own rows + halo rows already assembled)."""
    p2 = cfg.conv2
    y, c_base = cpu_ops.conv(pool1_slab, w.w2, w.b2, p2.F, p2.S, p2.P, H=Hp1, in_base=slab_base,
                             rows=rp.conv2_rows)
    cpu_ops.relu_(y)
    Hc2 = (Hp1 - p2.F + 2 * p2.P) // p2.S + 1
    y, _ = cpu_ops.maxpool(y, p2.F_pool, p2.S_pool, H=Hc2, in_base=c_base, rows=rp.out_rows)
    return cpu_ops.lrn(y, p2, cfg.lrn_mode)


def forward_torch_reference(x: torch.Tensor, w: Blocks12Weights, cfg: NetConfig,
                            upto: str = "lrn2") -> torch.Tensor:
    """Plain PyTorch fp32 restatement (tests only).  LRN is written out by hand because
    torch's local_response_norm hard-wires alpha/N; both conventions are needed."""
    import torch.nn.functional as Fn
    p1, p2 = cfg.conv1, cfg.conv2
    dev = x.device
    y = x.permute(0, 3, 1, 2).to(torch.float32)
    y = Fn.conv2d(y, w.w1.to(dev), w.b1.to(dev), stride=p1.S, padding=p1.P)
    if upto == "conv1":
        return y.permute(0, 2, 3, 1).contiguous()
    y = Fn.max_pool2d(Fn.relu(y), p1.F_pool, p1.S_pool)
    if upto == "pool1":
        return y.permute(0, 2, 3, 1).contiguous()
    y = Fn.conv2d(y, w.w2.to(dev), w.b2.to(dev), stride=p2.S, padding=p2.P)
    if upto == "conv2":
        return y.permute(0, 2, 3, 1).contiguous()
    y = Fn.max_pool2d(Fn.relu(y), p2.F_pool, p2.S_pool)
    if upto == "pool2":
        return y.permute(0, 2, 3, 1).contiguous()
    y = lrn_torch(y, p2.N_lrn, p2.alpha, p2.beta, p2.k_lrn, cfg.lrn_mode)
    return y.permute(0, 2, 3, 1).contiguous()


def lrn_torch(y_nchw: torch.Tensor, N: int, alpha: float, beta: float, k: float,
              mode: str) -> torch.Tensor:
    if N <= 0:
        return y_nchw
    half = N // 2
    sq = y_nchw * y_nchw
    Cc = y_nchw.shape[1]
    pad = torch.nn.functional.pad(sq, (0, 0, 0, 0, half, half))
    ss = sum(pad[:, i:i + Cc] for i in range(2 * half + 1))
    a = alpha / N if mode == "alpha_over_n" else alpha
    return y_nchw / (k + a * ss) ** beta
