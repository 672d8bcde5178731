"""Shape helpers.

``conv_out_dim`` / ``pool_out_dim`` follow
ref final_project/v2_mpi_only/2.2_scatter_halo/include/alexnet.hpp:35-44; the
``*_or_zero`` variants follow V4's, which return 0 when the window does not fit
(ref final_project/v4_mpi_cuda/include/alexnet.hpp:28-33).
"""
from __future__ import annotations

from dataclasses import dataclass

from .config import NetConfig


def conv_out_dim(D: int, F: int, P: int, S: int) -> int:
    if S <= 0:
        raise ValueError("stride must be positive")
    return (D - F + 2 * P) // S + 1


def pool_out_dim(D: int, F: int, S: int) -> int:
    if S <= 0:
        raise ValueError("stride must be positive")
    return (D - F) // S + 1


def conv_out_dim_or_zero(D: int, F: int, P: int, S: int) -> int:
    return 0 if D + 2 * P < F else conv_out_dim(D, F, P, S)


def pool_out_dim_or_zero(D: int, F: int, S: int) -> int:
    return 0 if D < F else pool_out_dim(D, F, S)


@dataclass(frozen=True)
class Block12Shapes:
    H: int; W: int; C: int
    Hc1: int; Wc1: int; K1: int
    Hp1: int; Wp1: int
    Hc2: int; Wc2: int; K2: int
    Hp2: int; Wp2: int

    @property
    def out_hwc(self):
        return (self.Hp2, self.Wp2, self.K2)

    @property
    def out_numel(self):
        return self.Hp2 * self.Wp2 * self.K2

    def flops_per_image(self) -> int:
        """2*MACs of conv1+conv2 (interior-tap count)."""
        c1 = self.Hc1 * self.Wc1 * self.K1 * self.C * 11 * 11
        c2 = self.Hc2 * self.Wc2 * self.K2 * self.K1 * 5 * 5
        return 2 * (c1 + c2)


def block12_shapes(cfg: NetConfig) -> Block12Shapes:
    p1, p2 = cfg.conv1, cfg.conv2
    Hc1 = conv_out_dim_or_zero(cfg.H, p1.F, p1.P, p1.S)
    Wc1 = conv_out_dim_or_zero(cfg.W, p1.F, p1.P, p1.S)
    Hp1 = pool_out_dim_or_zero(Hc1, p1.F_pool, p1.S_pool)
    Wp1 = pool_out_dim_or_zero(Wc1, p1.F_pool, p1.S_pool)
    Hc2 = conv_out_dim_or_zero(Hp1, p2.F, p2.P, p2.S)
    Wc2 = conv_out_dim_or_zero(Wp1, p2.F, p2.P, p2.S)
    Hp2 = pool_out_dim_or_zero(Hc2, p2.F_pool, p2.S_pool)
    Wp2 = pool_out_dim_or_zero(Wc2, p2.F_pool, p2.S_pool)
    return Block12Shapes(cfg.H, cfg.W, cfg.C, Hc1, Wc1, p1.K, Hp1, Wp1,
                         Hc2, Wc2, p2.K, Hp2, Wp2)
