#!/usr/bin/env python3
"""NCCL baseline: the same row decomposition and the same compute kernels, but
every transfer is a library collective instead of peer memory inside the kernels:

    scatter  : NCCL send/recv of each rank's input rows            (vs P2P pull fused into conv1)
    halo     : NCCL batch isend/irecv of pool1 rows                (vs peer loads in conv2's loader)
    gather   : NCCL send/recv of the final rows to rank 0          (vs P2P stores from the LRN epilogue)

"A path that only calls NCCL for halo, scatter or gather is the baseline, not the product"
(BASELINE.json).  Run under torchrun / bin/mpirun; prints device-timed images/s (max over ranks).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as td

import b200cnn
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem
from b200cnn.ops import cuda as cu
from b200cnn.ops._ctypes import LrnParams, LRN_MODES
from b200cnn.parallel import dist as bdist
from b200cnn.parallel.planner import make_plan


class NcclHaloRank:
    def __init__(self, cfg, w, ctx, precision="tf32", sharded=False):
        self.cfg, self.ctx, self.precision, self.sharded = cfg, ctx, precision, sharded
        self.plan = make_plan(cfg, ctx.world, "halo")
        self.rp = self.plan.ranks[ctx.rank]
        sh = self.sh = self.plan.shapes
        dev = ctx.device
        N, rp = cfg.batch, self.rp
        self.b1, self.b2 = w.b1.to(dev), w.b2.to(dev)
        if precision == "tf32":
            from b200cnn.ops import tc
            self.k1, self.k2 = tc.TcConv(w.w1, dev), tc.TcConv(w.w2, dev)
        else:
            self.k1, self.k2 = cu.pack_weights_ffck(w.w1).to(dev), cu.pack_weights_ffck(w.w2).to(dev)
        i0, i1 = rp.in_rows
        self.x = torch.zeros((N, max(i1 - i0, 1), sh.W, sh.C), device=dev)
        n0, n1 = rp.pool1_need
        self.p1 = torch.zeros((N, max(n1 - n0, 1), sh.Wp1, sh.K1), device=dev)
        a, b = rp.out_rows
        self.y = torch.zeros((N, max(b - a, 1), sh.Wp2, sh.K2), device=dev)
        self.full_in = None
        self.out = torch.zeros((N, sh.Hp2, sh.Wp2, sh.K2), device=dev) if ctx.rank == 0 else None
        self.status = torch.zeros(4, dtype=torch.int32, device=dev)

    def _block(self, k, bias, out_t, g, round_out=False):
        if self.precision == "tf32":
            k.run(bias, out_t.data_ptr(), g, round_out=round_out)
        else:
            cu.simt_block(k, bias, out_t.data_ptr(), g)

    def forward(self):
        cfg, sh, rp, ctx = self.cfg, self.sh, self.rp, self.ctx
        p1, p2 = cfg.conv1, cfg.conv2
        # ---- scatter (rank 0 -> all); skipped when the rows are already in place (--sharded) ------
        ops = []
        if self.sharded:
            pass
        elif ctx.rank == 0:
            for q in self.plan.ranks[1:]:
                if q.in_rows[1] > q.in_rows[0]:
                    ops.append(td.P2POp(td.isend, self.full_in[:, q.in_rows[0]:q.in_rows[1]].contiguous(), q.rank))
            self.x.copy_(self.full_in[:, rp.in_rows[0]:rp.in_rows[1]])
        elif rp.in_rows[1] > rp.in_rows[0]:
            ops.append(td.P2POp(td.irecv, self.x, 0))
        if ops:
            for r in td.batch_isend_irecv(ops):
                r.wait()
        # ---- block 1 into the "need" slab at the owned rows ---------------------------------------
        n0, n1 = rp.pool1_need
        o0, o1 = rp.pool1_own
        if o1 > o0:
            g = cu.make_block_geom(N=cfg.batch, H=sh.H, W=sh.W, Cin=sh.C, K=p1.K, F=p1.F, S=p1.S, P=p1.P,
                                   Hc=sh.Hc1, Wc=sh.Wc1, Hp=sh.Hp1, Wp=sh.Wp1, rows=(o0, o1), out_base=n0,
                                   out_slab=self.p1.shape[1], lrn=None,
                                   segs=[dict(ptr=self.x.data_ptr(), base=rp.in_rows[0], slab=self.x.shape[1],
                                              lo=rp.in_rows[0], hi=rp.in_rows[1])],
                                   status_ptr=self.status.data_ptr())
            self._block(self.k1, self.b1, self.p1, g, round_out=True)
        # ---- halo exchange (NCCL) ---------------------------------------------------------------
        ops, rbuf = [], []
        for peer, a, b in rp.halo:
            t = torch.empty((cfg.batch, b - a, sh.Wp1, sh.K1), device=ctx.device)
            rbuf.append((a, b, t))
            ops.append(td.P2POp(td.irecv, t, peer))
        for peer, a, b in rp.serves:
            ops.append(td.P2POp(td.isend, self.p1[:, a - n0:b - n0].contiguous(), peer))
        if ops:
            for r in td.batch_isend_irecv(ops):
                r.wait()
        for a, b, t in rbuf:
            self.p1[:, a - n0:b - n0].copy_(t)
        # ---- block 2 ----------------------------------------------------------------------------
        a, b = rp.out_rows
        if b > a:
            lrn = LrnParams(p2.N_lrn, p2.alpha, p2.beta, p2.k_lrn, LRN_MODES[cfg.lrn_mode])
            g = cu.make_block_geom(N=cfg.batch, H=sh.Hp1, W=sh.Wp1, Cin=sh.K1, K=p2.K, F=p2.F, S=p2.S, P=p2.P,
                                   Hc=sh.Hc2, Wc=sh.Wc2, Hp=sh.Hp2, Wp=sh.Wp2, rows=(a, b), out_base=a,
                                   out_slab=self.y.shape[1], lrn=lrn,
                                   segs=[dict(ptr=self.p1.data_ptr(), base=n0, slab=self.p1.shape[1], lo=n0, hi=n1)],
                                   status_ptr=self.status.data_ptr())
            self._block(self.k2, self.b2, self.y, g)
        # ---- gather (all -> rank 0) ---------------------------------------------------------------
        ops, rbuf = [], []
        if ctx.rank == 0:
            if b > a:
                self.out[:, a:b].copy_(self.y)
            for q in self.plan.ranks[1:]:
                if q.active:
                    t = torch.empty((cfg.batch, q.out_rows[1] - q.out_rows[0], sh.Wp2, sh.K2), device=ctx.device)
                    rbuf.append((q.out_rows, t))
                    ops.append(td.P2POp(td.irecv, t, q.rank))
        elif b > a:
            ops.append(td.P2POp(td.isend, self.y, 0))
        if ops:
            for r in td.batch_isend_irecv(ops):
                r.wait()
        for (qa, qb), t in rbuf:
            self.out[:, qa:qb].copy_(t)
        return self.out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--precision", default="tf32")
    ap.add_argument("--check", action="store_true", help="compare with the fused engine's output")
    ap.add_argument("--sharded", action="store_true",
                    help="every rank already holds its input rows (no scatter in the timed step): the like-for-like "
                         "baseline of bench.py's sharded placement; halo and gather still go through NCCL")
    a = ap.parse_args()
    ctx = bdist.init()
    cfg = b200cnn.default_config(batch=a.batch, init=INIT_RAND, seed=1, precision=a.precision)
    x, w = make_problem(cfg)
    r = NcclHaloRank(cfg, w, ctx, a.precision, a.sharded)
    if a.sharded:
        i0, i1 = r.rp.in_rows
        if i1 > i0:
            r.x.copy_(x[:, i0:i1].to(ctx.device))
    elif ctx.rank == 0:
        r.full_in = x.to(ctx.device)
    for _ in range(a.warmup):
        r.forward()
    torch.cuda.synchronize(); ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        out = r.forward()
    e1.record(); e1.synchronize()
    ms = ctx.max_float(e0.elapsed_time(e1) / a.steps)
    res = {"baseline": "nccl send/recv " + ("halo+gather (inputs sharded)" if a.sharded else "scatter+halo+gather") + ", unfused",
           "n_gpus": ctx.world, "batch": a.batch,
           "ms_per_step": round(ms, 4), "images_per_s": round(a.batch / ms * 1e3, 1), "precision": a.precision}
    if a.check and ctx.rank == 0 and ctx.world >= 1:
        from b200cnn.models.engine import single_gpu_engine
        ref = single_gpu_engine(cfg, w, ctx.device, a.precision).forward(x.to(ctx.device))
        res["max_abs_diff_vs_fused_np1"] = float((out - ref).abs().max())
    if ctx.rank == 0:
        print(json.dumps(res), flush=True)
    ctx.finalize()


if __name__ == "__main__":
    main()
