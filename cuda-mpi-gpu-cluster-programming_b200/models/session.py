"""Public inference API: host tensors in, host tensors out.

``Session`` is what a user of the reference's ``./template`` / ``mpirun -np N ./template``
switches to when they want the forward pass as a library call:

    sess = Session(cfg, weights)                  # one GPU
    y = sess.infer(x_pinned)                      # [N,H,W,C] fp32 host -> [N,13,13,256] host

    sess = Session(cfg, weights, dist=ctx)        # one process per GPU (torchrun / bin/mpirun)
    y = sess.infer(x_host_on_rank0)               # scatter="p2p": rank 0 owns the images
    y = sess.infer(my_row_slab)                   # scatter="sharded": each rank feeds its slab

Every call does the host->device copy of that call's input from pinned memory, the forward, and
the device->host copy of the result.  ``submit``/``result`` expose the same thing pipelined:
double-buffered input and output slots, a copy-in stream, a compute stream and a copy-out
stream, so step i+1's H2D overlaps step i's kernels and step i-1's D2H.  (The reference does
synchronous pageable cudaMemcpy inside its timed region,
ref v3_cuda_only/src/alexnet_cuda.cu:67-71,88.)
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from ..config import NetConfig
from ..parallel.symm import LocalSymm, IpcSymm
from .engine import Blocks12Engine
from .init import Blocks12Weights


class Session:
    """``gather``: where the result is assembled.  ``"rank0"`` = in rank 0's device tensor (every rank's LRN
    epilogue stores its rows there over NVLink) and rank 0 copies it to the host; ``"host"`` = every rank copies
    the rows it owns straight into ONE pinned host tensor shared by all ranks (parallel/hostshm.py), so the
    result leaves the node's GPUs over all their PCIe links at once.  ``row_groups``: see parallel/planner.py
    (None = the reference's pure row split, "auto" = two row groups x world/2 batch groups).  ``slots`` input /
    output slots are cycled; a result returned by :meth:`result` stays valid for ``slots - 2`` further submits."""

    def __init__(self, cfg: NetConfig, weights: Blocks12Weights, dist=None, device=None,
                 precision: Optional[str] = None, strategy: str = "halo", scatter: str = "p2p",
                 row_groups=None, gather: str = "rank0", slots: int = 2, halo: str = "push"):
        if gather not in ("rank0", "host", "local"):    # "local": rows stay on the rank that computed them (no D2H)
            raise ValueError(gather)
        self.cfg = cfg
        self.dist = dist
        # run on the CPUs next to this GPU before any pinned buffer is touched (see utils/numa.py)
        from ..utils.numa import bind_host_to_gpu
        self.host_binding = bind_host_to_gpu(device or (dist.device if dist else "cuda:0"))
        world = dist.world if dist is not None else 1
        rank = dist.rank if dist is not None else 0
        self.rank, self.world, self.slots = rank, world, slots
        if world > 1:
            self.symm = IpcSymm(dist)
        else:
            self.symm = LocalSymm(1, device or (dist.device if dist else "cuda:0")).view(0)
        self.gather = gather if world > 1 else "rank0"
        self.engine = Blocks12Engine(cfg, weights, self.symm, rank, world, precision=precision,
                                     strategy=strategy, scatter=scatter, row_groups=row_groups,
                                     gather="rank0" if self.gather == "rank0" else "local", slots=slots, halo=halo)
        self.symm.commit()
        self.engine.finalize()
        e = self.engine
        sh = e.sh
        self.device = e.device
        with torch.cuda.device(self.device):
            self.s_in = torch.cuda.Stream()
            self.s_compute = torch.cuda.Stream()
            self.s_out = torch.cuda.Stream()
            self.ev_h2d = [torch.cuda.Event() for _ in range(slots)]
            self.ev_compute = [torch.cuda.Event() for _ in range(slots)]
            self.ev_d2h = [torch.cuda.Event() for _ in range(slots)]
        self.step = 0
        self._slot_step = [0] * slots        # step number (1-based) last submitted into each slot
        self.h_out = None
        self.shared = None
        full = (cfg.batch, sh.Hp2, sh.Wp2, sh.K2)
        if self.gather == "host":
            from ..parallel.hostshm import SharedHostTensor
            self.shared = SharedHostTensor(full, slots, dist, self.device)
            if rank == 0:
                self.h_out = self.shared.views
            self.d2h_bytes = e.out_imgs * e.out_rows * sh.Wp2 * sh.K2 * 4
        elif self.gather == "local":
            self.d2h_bytes = 0
        else:
            if rank == 0:
                self.h_out = [torch.empty(full, dtype=torch.float32, pin_memory=True) for _ in range(slots)]
            self.d2h_bytes = (self.h_out[0].numel() * 4) if rank == 0 else 0
        self.h2d_bytes = e.in_bytes

    # -- what this rank must feed ---------------------------------------------------------------
    @property
    def input_rows(self):
        """Global input rows this rank supplies to ``infer``/``submit`` ((0,H) on rank 0 in p2p
        mode, nothing on other ranks; the rank's slab in sharded mode)."""
        e = self.engine
        return (e.in_base, e.in_base + e.in_rows)

    @property
    def input_images(self):
        """Images of the global batch this rank supplies (all of them on rank 0 in p2p mode; the share of
        the rank's batch group in sharded mode)."""
        e = self.engine
        if e.scatter == "p2p":
            return (0, self.cfg.batch) if self.rank == 0 else (0, 0)
        return e.rp.images

    def my_input(self, x: torch.Tensor) -> Optional[torch.Tensor]:
        """The part of a whole-batch host tensor [N,H,W,C] this rank feeds (None: nothing)."""
        (n0, n1), (lo, hi) = self.input_images, self.input_rows
        if n1 <= n0 or hi <= lo:
            return None
        return x[n0:n1, lo:hi]

    def submit(self, x_host: Optional[torch.Tensor]) -> int:
        e = self.engine
        i = self.step % self.slots
        self.step += 1
        self._slot_step[i] = self.step
        with torch.cuda.device(self.device):
            if e.in_rows > 0 and e.in_imgs > 0:
                assert x_host is not None and x_host.is_pinned(), "feed pinned host memory"
                # slot i is free once the forward that last read it has finished
                self.s_in.wait_event(self.ev_compute[i])
                with torch.cuda.stream(self.s_in):
                    e.h2d(i, x_host)
                    self.ev_h2d[i].record(self.s_in)
                self.s_compute.wait_event(self.ev_h2d[i])
            # output slot i must have been drained by its previous D2H
            self.s_compute.wait_event(self.ev_d2h[i])
            with torch.cuda.stream(self.s_compute):
                out = e.forward_enqueue(i, i)
                self.ev_compute[i].record(self.s_compute)
            if out is not None and self.gather != "local":
                self.s_out.wait_event(self.ev_compute[i])
                with torch.cuda.stream(self.s_out):
                    if self.shared is not None:      # my rows of my images -> their place in the shared tensor
                        sh = e.sh
                        row = sh.Wp2 * sh.K2 * 4
                        dst = (self.shared.slot_host_ptr(i)
                               + (e.rp.images[0] * sh.Hp2 + e.out_row0) * row)
                        from ..ops import cuda as cu
                        cu.memcpy2d_d2h(dst, sh.Hp2 * row, e.buf_out[i].ptr, e.out_rows * row,
                                        e.out_rows * row, e.out_imgs)
                        cu.host_flag_store(self.shared.flag_dev_ptr(i), self.step)
                    else:
                        self.h_out[i].copy_(out, non_blocking=True)
                    self.ev_d2h[i].record(self.s_out)
        return i

    def result(self, ticket: int) -> Optional[torch.Tensor]:
        """Blocks until the ticket's output is in host memory (rank 0) / this rank's part of it is done."""
        if self.shared is not None:
            self.ev_d2h[ticket].synchronize()
            if self.rank != 0:
                return None
            self.shared.wait_all(ticket, self._slot_step[ticket], self.engine._active)
            return self.h_out[ticket]
        if self.h_out is not None:
            self.ev_d2h[ticket].synchronize()
            return self.h_out[ticket]
        self.ev_compute[ticket].synchronize()
        return None

    def infer(self, x_host: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        if x_host is not None and not x_host.is_pinned():
            x_host = x_host.contiguous().pin_memory()
        y = self.result(self.submit(x_host))
        self.engine.check_status()
        return y

    def drain(self) -> None:
        with torch.cuda.device(self.device):
            for s in (self.s_in, self.s_compute, self.s_out):
                s.synchronize()
