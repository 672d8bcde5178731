"""Input / parameter initialisers.

Two modes, both from the reference:

* ``const`` — V2.1/V2.2/V3/V4: input 1.0, weights 0.01, biases 0.0
  (ref v4_mpi_cuda/src/main_mpi_cuda.cpp:29-33, v3_cuda_only/src/main_cuda.cpp:16-27).
* ``rand``  — V1: ``rand()/RAND_MAX*0.1`` data, ``(rand()/RAND_MAX-0.5)*0.02`` weights,
  biases 0.1, drawn in the order input, W1, W2 (ref v1_serial/src/alexnet_serial.cpp:39-57,
  main.cpp:51-62).  V1 seeds with ``time(0)``; here the seed is explicit and the glibc
  generator is reproduced natively, so a seed pins the stream.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch

from ..config import NetConfig, INIT_CONST, INIT_RAND


@dataclass
class Blocks12Weights:
    w1: torch.Tensor   # [K1, C, F1, F1]   (reference layout [K][C][F][F])
    b1: torch.Tensor   # [K1]
    w2: torch.Tensor   # [K2, K1, F2, F2]
    b2: torch.Tensor   # [K2]

    def to(self, device) -> "Blocks12Weights":
        return Blocks12Weights(*(t.to(device) for t in (self.w1, self.b1, self.w2, self.b2)))

    def state_dict(self):
        return {"w1": self.w1, "b1": self.b1, "w2": self.w2, "b2": self.b2}


def make_problem(cfg: NetConfig):
    """Returns (input [N,H,W,C] fp32 CPU, Blocks12Weights on CPU)."""
    p1, p2 = cfg.conv1, cfg.conv2
    n_in = cfg.batch * cfg.H * cfg.W * cfg.C
    s1 = (p1.K, cfg.C, p1.F, p1.F)
    s2 = (p2.K, p1.K, p2.F, p2.F)
    if cfg.init == INIT_CONST:
        x = torch.full((cfg.batch, cfg.H, cfg.W, cfg.C), 1.0, dtype=torch.float32)
        w = Blocks12Weights(torch.full(s1, 0.01), torch.zeros(p1.K),
                            torch.full(s2, 0.01), torch.zeros(p2.K))
        return x, w
    if cfg.init == INIT_RAND:
        from ..ops.cpu import GlibcRand
        g = GlibcRand(cfg.seed)
        x = g.data(n_in).view(cfg.batch, cfg.H, cfg.W, cfg.C)
        w1 = g.weights(s1[0] * s1[1] * s1[2] * s1[3]).view(s1)
        w2 = g.weights(s2[0] * s2[1] * s2[2] * s2[3]).view(s2)
        w = Blocks12Weights(w1, torch.full((p1.K,), 0.1), w2, torch.full((p2.K,), 0.1))
        return x, w
    raise ValueError(cfg.init)
