"""GPU engine for AlexNet blocks 1-2: resident weights and activations, two fused kernels per
forward, peer-memory scatter / halo / gather.

One :class:`Blocks12Engine` is one rank.  It is the replacement for the reference's L3
orchestrators — ``alexnetForwardPassCUDA`` (10 cudaMalloc + 5 H2D + 7 launches + D2H + 10
cudaFree per call, ref v3_cuda_only/src/alexnet_cuda.cu:22-95), ``alexnetTileForwardCUDA``
(re-uploads 2.6 MB of weights per call, ref v4_mpi_cuda/src/alexnet_mpi_cuda.cu:157-205) and the
decomposition logic inlined in V4's main (ref v4_mpi_cuda/src/main_mpi_cuda.cpp:52-130):

* buffers are allocated once (symmetric memory), weights are packed and uploaded once;
* block 1 = conv1+bias+ReLU+pool1 in ONE kernel whose loader pulls the input rows it needs from
  rank 0's image buffer over NVLink (the scatter) and whose epilogue bumps per-image arrival
  counters on the ranks that will read its pool1 rows;
* block 2 = conv2+bias+ReLU+pool2+LRN in ONE kernel whose loader reads halo pool1 rows straight
  from the owning peer (gated by those counters) and whose epilogue stores the final rows straight
  into rank 0's output tensor (the gather);
* no barrier sits on the forward: the pool1 slabs are double-buffered by forward parity, rank 0 opens
  forward k by bumping a ``go`` counter on every rank (one tiny launch), block 2 of forward k is gated
  on ``go >= k`` (off the critical path: it was published one conv1 earlier), and the last CTA of every
  rank's block 2 bumps a completion counter on rank 0 — so the whole multi-GPU forward is
  stream-ordered with no host synchronisation and no all-to-all (see ``_enqueue_stages``);
* the world is R row groups x B batch groups (``row_groups``): a rank splits rows only with the R-1
  ranks that hold the other rows of ITS images.

``precision``: ``"simt"`` exact FP32 (validation twin) or ``"tf32"`` tcgen05 tensor cores.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import torch

from ..config import NetConfig
from ..ops import cuda as cu
from ..ops._ctypes import LrnParams, LRN_MODES
from ..parallel.planner import make_plan, Plan, RankPlan
from ..parallel.symm import LocalSymm
from .init import Blocks12Weights


@dataclass
class _Layout:
    """int32 flag pad layout inside the per-rank "flags" buffer."""
    world: int
    nmax: int

    @property
    def arr_off(self):      # arrival counters [src_rank][image]
        return 0

    @property
    def bar_off(self):      # barrier counters [src_rank]
        return self.world * self.nmax

    @property
    def done_off(self):     # completion counters [src_rank]
        return self.bar_off + self.world

    @property
    def status_off(self):
        return self.done_off + self.world

    @property
    def sync_off(self):     # second barrier counter set [src_rank] for sync_ranks() (no epoch)
        return self.status_off + 2

    @property
    def epoch_off(self):    # device-side forward counter (bumped by the kernel that opens a forward)
        return self.status_off + 1

    @property
    def go_off(self):       # forwards opened by rank 0 (gate of block 2)
        return self.sync_off + self.world

    @property
    def cta_off(self):      # CTAs of block 2 that have finished, all forwards (completion notice)
        return self.go_off + 1

    @property
    def count(self):
        return self.cta_off + 1 + 16


class Blocks12Engine:
    def __init__(self, cfg: NetConfig, weights: Blocks12Weights, symm, rank: int = 0,
                 world: int = 1, *, precision: Optional[str] = None, strategy: str = "halo",
                 scatter: str = "p2p", timeout_ns: int = cu.DEFAULT_TIMEOUT_NS,
                 device_epoch: Optional[bool] = None, use_graph: Optional[bool] = None,
                 row_groups=None, gather: str = "rank0", slots: int = 2, halo: str = "push"):
        if scatter not in ("p2p", "sharded"):
            raise ValueError(scatter)
        if gather not in ("rank0", "local"):
            raise ValueError(gather)
        if halo not in ("push", "pull"):
            raise ValueError(halo)
        # halo="push": conv1's epilogue also stores the boundary pool1 rows into the neighbour's slab (P2P stores
        # over NVLink, then the arrival counter), so conv2 reads local memory only and can be fed by TMA;
        # halo="pull": conv2's loader reads them from the neighbour's slab (P2P loads in the producer warps).
        self.halo = halo
        # gather="rank0": the LRN epilogue stores every rank's rows into rank 0's output tensor over NVLink
        # (the reference's MPI_Gatherv); "local": every rank keeps the rows of the images it owns (callers
        # that want the result in HOST memory read each rank's rows back over that rank's own PCIe link).
        self.gather = gather
        self.cfg = cfg
        self.rank, self.world = rank, world
        self.symm = symm
        self.device = torch.device(symm.device)
        self.precision = precision or cfg.precision
        self.scatter = scatter
        self.timeout_ns = timeout_ns
        self.plan: Plan = make_plan(cfg, world, strategy, row_groups)
        self.rp: RankPlan = self.plan.ranks[rank]
        self.n_img, self.img0 = self.rp.n_images, self.rp.images[0]
        self.sh = self.plan.shapes
        self.epoch = 0
        self.launches = 0            # kernels launched by this engine (bench reports it)
        self.s2d = False
        # device_epoch: flag targets are derived on the device from a forward counter, so launch
        # arguments never change and a whole forward can be replayed as one CUDA graph.  The
        # single-GPU emulation of several ranks launches barrier halves separately and keeps the
        # host-side epoch instead.
        self.device_epoch = (world > 1) if device_epoch is None else device_epoch
        self.use_graph = True if use_graph is None else use_graph
        self._stage = None           # contiguous H2D landing buffers (pitched inputs only)
        self._graphs = {}            # (in_idx, out_idx, parity) -> (torch.cuda.CUDAGraph, kernels per replay)
        self._calls = {}
        self.lay = _Layout(world, cfg.batch)
        sh, rp = self.sh, self.rp
        N = self.n_img               # images this rank works on

        # ---- weights: packed once, resident ------------------------------------------------
        with torch.cuda.device(self.device):
            self.b1 = weights.b1.to(self.device).contiguous()
            self.b2 = weights.b2.to(self.device).contiguous()
            if self.precision == "simt":
                self.w1p = cu.pack_weights_ffck(weights.w1).to(self.device)
                self.w2p = cu.pack_weights_ffck(weights.w2).to(self.device)
            elif self.precision == "tf32":
                import os
                from ..ops import tc, ext
                p1 = cfg.conv1
                # stride-S conv1 on C=3 -> stride-1 conv on the space-to-depth tensor (S*S*C channels):
                # its im2col rows are aligned 16 B pieces of the tensor, which the loader cp.async's
                # straight into the MMA stage (no shared-memory patch, no smem->smem expansion).
                # (kept as an option, B200_S2D=1: the pitched-row direct loader below needs no extra pass)
                self.s2d = (p1.S > 1 and p1.P == 0 and (p1.S * p1.S * cfg.C) % 4 == 0
                            and bool(os.environ.get("B200_S2D")))
                self.tc1 = tc.TcConv(ext.s2d_weights(weights.w1, p1.S) if self.s2d else weights.w1, self.device)
                self.tc2 = tc.TcConv(weights.w2, self.device)
            else:
                raise ValueError(self.precision)

        # ---- symmetric buffers ---------------------------------------------------------------
        f4 = 4
        self.n_in_buf = slots          # input / output slots a pipelined caller cycles through
        if scatter == "p2p":                 # rank 0 holds every image of the batch, whole
            in_rows = sh.H if rank == 0 else 0
            self.in_base = 0
            self.in_imgs = cfg.batch if rank == 0 else 0
        else:
            in_rows = rp.in_rows[1] - rp.in_rows[0] if N > 0 else 0
            self.in_base = rp.in_rows[0]
            self.in_imgs = N
        self.in_rows = in_rows
        # Row pitch of the resident input: padded to a multiple of 4 floats (and long enough for the last
        # pixel's filter-row run) so conv1's loader can cp.async 16 B chunks straight from it; the H2D
        # copy is a pitched DMA (cudaMemcpy2DAsync), so the padding costs nothing.
        p1 = cfg.conv1
        self.in_pitch = sh.W * sh.C
        if self.precision == "tf32" and not self.s2d and p1.P == 0 and (p1.S * sh.C) % 4 == 0:
            run = (p1.F * sh.C + 3) // 4 * 4
            self.in_pitch = (max(sh.W * sh.C, (sh.Wc1 - 1) * p1.S * sh.C + run) + 3) // 4 * 4
        self.in_bytes = self.in_imgs * in_rows * sh.W * sh.C * f4   # payload bytes per step (H2D accounting)
        self.in_alloc = self.in_imgs * in_rows * self.in_pitch * f4 + 256
        self.buf_in = [symm.alloc(f"in{i}", self.in_alloc) for i in range(self.n_in_buf)]
        if self.s2d:
            S1 = cfg.conv1.S
            self.s2d_ws = -(-sh.W // S1)
            self.s2d_hs = -(-sh.H // S1)
            self.s2d_c = S1 * S1 * sh.C
            if in_rows > 0:
                self.s2d_y0 = self.in_base // S1
                self.s2d_ys = -(-(self.in_base + in_rows) // S1) - self.s2d_y0
            else:
                self.s2d_y0, self.s2d_ys = 0, 0
            self.buf_s2d = [symm.alloc(f"s2d{i}", self.in_imgs * self.s2d_ys * self.s2d_ws * self.s2d_c * f4)
                            for i in range(self.n_in_buf)]
        self.p1_base, self.p1_rows = self._pool1_slab(rp)
        own = self.p1_rows
        # two pool1 slabs, used alternately by forward parity: a neighbour may still read forward k's rows
        # while this rank's conv1 already writes forward k+1's
        self.buf_pool1 = [symm.alloc(f"pool1_{i}", N * own * sh.Wp1 * sh.K1 * f4) for i in range(2)]
        self.n_out_buf = slots
        if gather == "rank0":
            self.out_imgs, self.out_row0, self.out_rows = (cfg.batch, 0, sh.Hp2) if rank == 0 else (0, 0, 0)
        else:
            self.out_imgs, self.out_row0 = N, rp.out_rows[0]
            self.out_rows = rp.out_rows[1] - rp.out_rows[0]
        self.buf_out = [symm.alloc(f"out{i}", self.out_imgs * self.out_rows * sh.Wp2 * sh.K2 * f4)
                        for i in range(self.n_out_buf)]
        self.buf_flags = symm.alloc("flags", self.lay.count * 4)
        self._final = False

    # ------------------------------------------------------------------------------------------
    def finalize(self) -> None:
        """Call after every rank allocated and ``symm.commit()`` ran: resolves peer pointers."""
        self._final = True
        sh = self.sh
        if self.out_imgs * self.out_rows > 0:
            self.outs = [b.tensor((self.out_imgs, self.out_rows, sh.Wp2, sh.K2)) for b in self.buf_out]
        else:
            self.outs = [None] * self.n_out_buf
        self.out = self.outs[0]
        self._active = [r for r in range(self.world) if self.plan.ranks[r].active]
        self.flags_t = self.buf_flags.tensor((self.lay.count,), torch.int32)
        # arrival-counter targets: tiles per image of the peer's block 1 that produce the rows read from it
        self._b1_tiles = {}
        for peer, lo, hi in self.rp.halo:
            self._b1_tiles[peer] = self._tiles_rows(self._geom_block1(peer, probe=True), lo, hi)

    def _pool1_slab(self, rp: RankPlan):
        """(first global row, rows) of a rank's pool1 slab: the rows it owns, plus — when halos are pushed — the
        neighbours' rows its conv2 reads."""
        lo, hi = rp.pool1_own
        if self.halo == "push" and rp.halo:
            lo, hi = min(lo, rp.pool1_need[0]), max(hi, rp.pool1_need[1])
        return lo, max(hi - lo, 0)

    def _flag_ptr(self, rank: int, int_index: int) -> int:
        return self.symm.ptr(rank, "flags") + 4 * int_index

    def _tiles_rows(self, g, lo: int, hi: int) -> int:
        if self.precision == "simt":
            return cu.simt_block_tiles_rows(g, lo, hi)
        from ..ops import tc
        return tc.block_tiles_rows(g, lo, hi)

    # ---- geometry ------------------------------------------------------------------------------
    def input_view(self, i: int = 0) -> torch.Tensor:
        """This rank's input buffer ``i`` as an [N, rows, W, C] view (rows are ``in_pitch`` floats apart;
        rank 0 holds the whole batch in p2p-scatter mode)."""
        sh = self.sh
        n, rows, pitch = self.in_imgs, self.in_rows, self.in_pitch
        flat = self.buf_in[i].tensor((max(n * rows * pitch, 1),))
        return flat.as_strided((n, rows, sh.W, sh.C), (rows * pitch, pitch, sh.C, 1))

    def pool1_view(self) -> torch.Tensor:
        """This rank's pool1 rows of the latest forward, [n_img, own rows, Wp1, K1]."""
        sh, (a, b) = self.sh, self.rp.pool1_own
        t = self.buf_pool1[self.epoch & 1].tensor((self.n_img, self.p1_rows, sh.Wp1, sh.K1))
        return t[:, a - self.p1_base:b - self.p1_base]

    def h2d(self, i: int, x_host: torch.Tensor) -> None:
        """Pinned host [N, rows, W, C] -> input buffer ``i`` on the current stream.  A pitched input is
        filled in two steps — one contiguous DMA at full PCIe rate into a staging buffer, then a ~30 us
        device copy that spreads the rows to their pitch (a 2-D DMA of 2.7 KB rows is 3x slower)."""
        sh = self.sh
        assert x_host.is_contiguous() and tuple(x_host.shape) == (self.in_imgs, self.in_rows, sh.W, sh.C)
        with torch.cuda.device(self.device):
            if self.in_pitch == sh.W * sh.C:
                self.input_view(i).copy_(x_host, non_blocking=True)
                return
            if self._stage is None:
                self._stage = [torch.empty(x_host.shape, dtype=torch.float32, device=self.device)
                               for _ in range(self.n_in_buf)]
            self._stage[i].copy_(x_host, non_blocking=True)
            from ..ops import ext
            ext.repitch_rows_(self._stage[i].data_ptr(), self.buf_in[i].ptr, self.in_imgs * self.in_rows,
                              sh.W * sh.C, self.in_pitch)

    def _geom_block1(self, r: int, probe: bool = False, in_idx: int = 0):
        cfg, sh, p1 = self.cfg, self.sh, self.cfg.conv1
        rp = self.plan.ranks[r]
        img_bytes = rp.images[0] * 4                      # x floats per image: offset into rank 0's batch
        s2d = self.s2d
        S1 = p1.S
        Hin = self.s2d_hs if s2d else sh.H
        mirrors, signal_rows = [], []
        if probe:
            segs = [dict(ptr=1, base=0, slab=Hin, lo=0, hi=Hin)]
            signals, out_ptr = [], 0
        else:
            if self.scatter == "p2p":                  # this rank's images inside rank 0's whole batch
                name = f"s2d{in_idx}" if s2d else f"in{in_idx}"
                per_img = Hin * (self.s2d_ws * self.s2d_c if s2d else self.in_pitch)
                segs = [dict(ptr=self.symm.ptr(0, name) + img_bytes * per_img, base=0, slab=Hin, lo=0, hi=Hin)]
            elif s2d:
                segs = [dict(ptr=self.buf_s2d[in_idx].ptr, base=self.s2d_y0, slab=self.s2d_ys,
                             lo=self.s2d_y0, hi=self.s2d_y0 + self.s2d_ys)]
            else:
                segs = [dict(ptr=self.buf_in[in_idx].ptr, base=self.in_base, slab=self.in_rows,
                             lo=rp.in_rows[0], hi=rp.in_rows[1])]
            # one arrival counter per reader, bumped only by the tiles that produce the rows it reads
            served = sorted(rp.serves)
            signals = [self._flag_ptr(peer, self.lay.arr_off + r * self.lay.nmax) for peer, _, _ in served]
            signal_rows = [(lo, hi) for _, lo, hi in served]
            if self.halo == "push":
                parity = self.epoch & 1
                for peer, lo, hi in rp.serves:
                    pb, prows = self._pool1_slab(self.plan.ranks[peer])
                    mirrors.append(dict(ptr=self.symm.ptr(peer, f"pool1_{parity}"), base=pb, slab=prows, lo=lo, hi=hi))
        # p2p scatter: the images are rank 0's upload of THIS forward, so conv1 on the other ranks may only start
        # once rank 0 has opened the forward (its stage_open follows the upload in stream order)
        gate = {}
        if not probe and self.scatter == "p2p" and self.world > 1 and r != 0:
            gate = dict(gate_ptr=self._flag_ptr(r, self.lay.go_off),
                        gate_target=1 if self.device_epoch else self.epoch,
                        epoch_ptr=self._flag_ptr(r, self.lay.epoch_off) if self.device_epoch else 0)
        conv = dict(H=Hin, W=self.s2d_ws, Cin=self.s2d_c, F=-(-p1.F // S1), S=1, P=0) if s2d else \
            dict(H=sh.H, W=sh.W, Cin=sh.C, F=p1.F, S=p1.S, P=p1.P, in_pitch=self.in_pitch)
        g = cu.make_block_geom(
            N=rp.n_images, K=p1.K, **conv,
            Hc=sh.Hc1, Wc=sh.Wc1, Hp=sh.Hp1, Wp=sh.Wp1, rows=rp.pool1_own,
            out_base=self._pool1_slab(rp)[0], out_slab=self._pool1_slab(rp)[1], lrn=None,
            segs=segs, signals=signals, signal_rows=signal_rows, mirrors=mirrors,
            status_ptr=0 if probe else self._flag_ptr(self.rank, self.lay.status_off),
            timeout_ns=self.timeout_ns, **gate)
        return g

    def _geom_block2(self, parity: int = 0):
        cfg, sh, p2, rp = self.cfg, self.sh, self.cfg.conv2, self.rp
        own0, own1 = rp.pool1_own
        segs = []
        if own1 > own0:
            segs.append(dict(ptr=self.buf_pool1[parity].ptr, base=self.p1_base, slab=self.p1_rows, lo=own0, hi=own1))
        for peer, lo, hi in rp.halo:
            if self.halo == "push":      # the neighbour's epilogue stored these rows into OUR slab
                src = dict(ptr=self.buf_pool1[parity].ptr, base=self.p1_base, slab=self.p1_rows)
            else:
                pb, prows = self._pool1_slab(self.plan.ranks[peer])
                src = dict(ptr=self.symm.ptr(peer, f"pool1_{parity}"), base=pb, slab=prows)
            segs.append(dict(**src,
                             lo=lo, hi=hi,
                             flag=self._flag_ptr(self.rank, self.lay.arr_off + peer * self.lay.nmax),
                             target=(1 if self.device_epoch else self.epoch) * self._b1_tiles[peer]))
        lrn = LrnParams(p2.N_lrn, p2.alpha, p2.beta, p2.k_lrn, LRN_MODES[cfg.lrn_mode])
        multi = self.world > 1
        per_fwd = 1 if self.device_epoch else self.epoch
        if self.gather == "rank0":
            out_base, out_slab = 0, sh.Hp2
        else:
            out_base, out_slab = self.out_row0, self.out_rows
        return cu.make_block_geom(
            N=self.n_img, H=sh.Hp1, W=sh.Wp1, Cin=sh.K1, K=p2.K, F=p2.F, S=p2.S, P=p2.P,
            Hc=sh.Hc2, Wc=sh.Wc2, Hp=sh.Hp2, Wp=sh.Wp2, rows=rp.out_rows,
            out_base=out_base, out_slab=out_slab, lrn=lrn, segs=segs, signals=[],
            status_ptr=self._flag_ptr(self.rank, self.lay.status_off), timeout_ns=self.timeout_ns,
            epoch_ptr=self._flag_ptr(self.rank, self.lay.epoch_off) if self.device_epoch else 0,
            # flow control: rank 0 has opened this forward, i.e. everything of the forward before the previous
            # one is finished everywhere (its output slot and the pool1 slabs of this parity are free)
            gate_ptr=self._flag_ptr(self.rank, self.lay.go_off) if multi and self.rank != 0 else 0,
            gate_target=per_fwd,
            done_count_ptr=self._flag_ptr(self.rank, self.lay.cta_off),
            done=[self._flag_ptr(0, self.lay.done_off + self.rank)] if multi else [])

    # ---- forward stages (each enqueues on the current stream) ----------------------------------
    def stage_begin(self) -> None:
        self.epoch += 1

    def stage_open(self) -> None:
        """Opens the forward on the device: bumps this rank's forward counter and, on rank 0, the ``go``
        counter of every other active rank.  Replaces the all-to-all barrier that used to open a forward
        (the reference's MPI_Barrier, ref v4_mpi_cuda/src/main_mpi_cuda.cpp:151): rank 0 gets here only
        after its completion wait of the previous forward, so ``go >= k`` tells a rank that every block 2
        of forward k-1 has finished everywhere."""
        if self.world > 1:
            go = [self._flag_ptr(r, self.lay.go_off) for r in self._active if r != 0] if self.rank == 0 else []
            if go or self.device_epoch:
                cu.flag_begin(self._flag_ptr(self.rank, self.lay.epoch_off) if self.device_epoch else 0, go)
                self.launches += 1

    def sync_ranks(self) -> None:
        """Stream-ordered device-side barrier on its own counters (does not open a forward): used to
        line the ranks up, e.g. right before a timed region.  One GPU per rank only."""
        if self.world > 1:
            self._sync_epoch = getattr(self, "_sync_epoch", 0) + 1
            W = range(self.world)
            with torch.cuda.device(self.device):
                cu.flag_barrier([self._flag_ptr(r, self.lay.sync_off + self.rank) for r in W],
                                [self._flag_ptr(self.rank, self.lay.sync_off + r) for r in W],
                                [self._sync_epoch] * self.world,
                                self._flag_ptr(self.rank, self.lay.status_off), self.timeout_ns, 0)

    def stage_s2d(self, in_idx: int = 0) -> None:
        """Space-to-depth of this rank's input slab (rank 0's whole batch in p2p-scatter mode).  Runs
        before the opening barrier, so peers only ever read a finished tensor."""
        if self.s2d and self.in_rows > 0:
            from ..ops import ext
            sh = self.sh
            assert self.in_pitch == sh.W * sh.C, "the s2d pre-pass reads an unpadded input"
            ext.space_to_depth_(self.buf_in[in_idx].ptr, self.buf_s2d[in_idx].ptr, self.in_imgs, sh.H, sh.W,
                                sh.C, self.cfg.conv1.S, self.in_base, self.in_rows, self.s2d_y0, self.s2d_ys)
            self.launches += 1

    def stage_block1(self, in_idx: int = 0) -> None:
        rp = self.rp
        if rp.pool1_own[1] <= rp.pool1_own[0]:
            return
        g = self._geom_block1(self.rank, in_idx=in_idx)
        pool1 = self.buf_pool1[self.epoch & 1].ptr
        if self.precision == "simt":
            cu.simt_block(self.w1p, self.b1, pool1, g)
        else:
            self.tc1.run(self.b1, pool1, g, round_out=True)
        self.launches += 1

    def stage_block2(self, out_idx: int = 0) -> None:
        if not self.rp.active:
            return
        g = self._geom_block2(self.epoch & 1)
        if self.gather == "rank0":       # this rank's images inside rank 0's [batch, Hp2, Wp2, K2] tensor
            sh = self.sh
            out_ptr = self.symm.ptr(0, f"out{out_idx}") + self.img0 * sh.Hp2 * sh.Wp2 * sh.K2 * 4
        else:
            out_ptr = self.buf_out[out_idx].ptr
        if self.precision == "simt":
            cu.simt_block(self.w2p, self.b2, out_ptr, g)
        else:
            self.tc2.run(self.b2, out_ptr, g, overlap_prev=True)     # prologue overlaps block 1's tail
        self.launches += 1

    def stage_done_wait(self) -> None:
        """Rank 0: every active rank's block 2 has stored its rows (their last CTAs bumped the counters)."""
        if self.world > 1 and self.rank == 0:
            cu.flag_wait([self._flag_ptr(0, self.lay.done_off + r) for r in self._active],
                         [1 if self.device_epoch else self.epoch] * len(self._active),
                         self._flag_ptr(0, self.lay.status_off), self.timeout_ns,
                         self._flag_ptr(0, self.lay.epoch_off) if self.device_epoch else 0)
            self.launches += 1

    def forward_enqueue(self, in_idx: int = 0, out_idx: int = 0) -> Optional[torch.Tensor]:
        """Enqueue one forward on the current stream (real multi-rank or single GPU).  The input
        must already be in ``input_view(in_idx)`` (rank 0 in p2p mode, every rank in sharded
        mode).  Returns rank 0's output tensor ``out_idx`` (valid once the stream gets here).
        Every rank must pass the same ``in_idx`` / ``out_idx`` for the same forward."""
        assert self._final, "call finalize() first"
        key = (in_idx, out_idx, (self.epoch + 1) & 1)     # launch arguments depend on the pool1 slab parity
        with torch.cuda.device(self.device):
            hit = self._graphs.get(key)
            if hit is not None:
                hit[0].replay()
                self.epoch += 1
                self.launches += hit[1]
                return self.outs[out_idx]
            n = self._calls.get(key, 0)
            self._calls[key] = n + 1
            graphable = self.use_graph and (self.world == 1 or self.device_epoch)
            if graphable and n >= 1:                      # first call ran eagerly (warm-up); capture now
                try:
                    l0 = self.launches
                    e0 = self.epoch
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        self._enqueue_stages(in_idx, out_idx)
                    self._graphs[key] = (g, self.launches - l0)
                    self.launches, self.epoch = l0, e0    # capture launched nothing
                    return self.forward_enqueue(in_idx, out_idx)
                except Exception as ex:
                    # never fall back silently: a forward that is timed as one graph replay must BE one (an eager
                    # multi-GPU forward is launch-bound and would be reported as if it were the product)
                    import os
                    if os.environ.get("B200_ALLOW_EAGER") != "1":
                        raise RuntimeError(f"[Rank {self.rank}] CUDA graph capture of the forward failed: {ex} "
                                           f"(use_graph=False or B200_ALLOW_EAGER=1 to run eagerly on purpose)") from ex
                    import warnings
                    warnings.warn(f"CUDA graph capture failed ({ex}); running eagerly (B200_ALLOW_EAGER=1)")
                    self.use_graph = False
            self._enqueue_stages(in_idx, out_idx)
        return self.outs[out_idx]

    def _enqueue_stages(self, in_idx: int, out_idx: int) -> None:
        """Forward k on one rank, stream-ordered:  open -> block 1 -> block 2 [-> completion wait on rank 0].
        Why no barrier is needed (k = forward number, slabs/slots indexed by k % 2):
        * block 1 of k overwrites the pool1 slab last read by the neighbours' block 2 of k-2.  It follows
          this rank's block 2 of k-1 in stream order, which was gated on go >= k-1, which rank 0 published
          after its completion wait of k-2: every block 2 of k-2 has finished.
        * block 2 of k writes rank 0's output slot last used by forward k-2 (the caller owns the k-2 vs k
          hand-over on rank 0; go >= k says rank 0 has begun k).
        * arrival counters only grow; a neighbour that runs ahead can only over-satisfy a wait."""
        self.stage_begin()
        self.stage_s2d(in_idx)
        self.stage_open()
        self.stage_block1(in_idx)
        self.stage_block2(out_idx)
        self.stage_done_wait()

    def forward(self, x: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
        """Device-tensor API: copies ``x`` (this rank's images and rows; the whole batch on rank 0 in p2p
        mode) into the resident buffer and runs one forward."""
        if x is not None and self.in_rows > 0:
            self.input_view(0).copy_(x, non_blocking=True)
        return self.forward_enqueue(0)

    def status(self) -> int:
        """First error latched by a device-side wait (0 = ok).  Synchronises."""
        torch.cuda.synchronize(self.device)
        return int(self.flags_t[self.lay.status_off].item())

    def check_status(self) -> None:
        s = self.status()
        if s != 0:
            raise RuntimeError(f"[Rank {self.rank}] device-side failure: status {s} "
                               f"(-4 = cross-GPU flag wait timed out)")


def single_gpu_engine(cfg: NetConfig, weights: Blocks12Weights, device="cuda:0",
                      precision: Optional[str] = None) -> Blocks12Engine:
    """V3-equivalent: one GPU, whole image, two kernels per forward."""
    symm = LocalSymm(1, device).view(0)
    e = Blocks12Engine(cfg, weights, symm, 0, 1, precision=precision)
    e.finalize()
    return e


class EmulatedWorld:
    """All ranks of a row-decomposed forward on ONE device in ONE process.  Kernels of different
    ranks are launched layer by layer, so every flag a kernel waits on is already set (spinning
    co-scheduled ranks on one GPU is forbidden).  Exercises the exact peer-pointer, halo, scatter
    and gather index math of the multi-GPU path."""

    def __init__(self, cfg: NetConfig, weights: Blocks12Weights, world: int, device="cuda:0",
                 precision: Optional[str] = None, strategy: str = "halo", scatter: str = "p2p",
                 row_groups=None, gather: str = "rank0", slots: int = 2, halo: str = "push"):
        self.symm = LocalSymm(world, device)
        self.engines: List[Blocks12Engine] = [
            Blocks12Engine(cfg, weights, self.symm.view(r), r, world, precision=precision,
                           strategy=strategy, scatter=scatter, device_epoch=False, use_graph=False,
                           row_groups=row_groups, gather=gather, halo=halo)
            for r in range(world)]
        for e in self.engines:
            e.finalize()
        self.cfg = cfg

    def load_input(self, x: torch.Tensor) -> None:
        for e in self.engines:
            if e.in_rows > 0:
                if e.scatter == "p2p":
                    e.input_view(0).copy_(x)
                else:
                    a, b = e.rp.in_rows
                    n0, n1 = e.rp.images
                    e.input_view(0).copy_(x[n0:n1, a:b])

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """One forward, layer by layer over the ranks (rank 0 first: its ``go`` is out before any gated
        launch runs).  With gather="local" the ranks' row blocks are assembled here."""
        self.load_input(x)
        E = self.engines
        for e in E: e.stage_begin()
        for e in E: e.stage_s2d(0)
        for e in E: e.stage_open()
        for e in E: e.stage_block1(0)
        for e in E: e.stage_block2()
        for e in E: e.stage_done_wait()
        for e in E: e.check_status()
        if E[0].gather == "rank0":
            return E[0].out
        sh = E[0].sh
        y = torch.empty((self.cfg.batch, sh.Hp2, sh.Wp2, sh.K2), dtype=torch.float32, device=E[0].device)
        for e in E:
            if e.out is not None:
                n0, n1 = e.rp.images
                y[n0:n1, e.out_row0:e.out_row0 + e.out_rows] = e.out
        return y
