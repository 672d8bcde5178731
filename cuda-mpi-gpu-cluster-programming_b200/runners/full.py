"""Full AlexNet forward (extension; BASELINE.json config 5), one process per GPU.

    mpirun -np N python -m b200cnn.runners.full --batch 256

Blocks 1-2 run row-decomposed across the N GPUs with the fused peer-memory kernels (exactly as V4/V5).
Conv3-5 + Pool5 work on 13x13 maps, too small to split by rows, so they are sharded by BATCH: after a
device-side barrier every rank copies its images of the gathered LRN2 tensor out of rank 0's buffer over
NVLink (one pass; the implicit-GEMM loader would otherwise re-read each pixel nine times over the link),
runs the three implicit-GEMM convolutions and the pool locally and stores its feature rows straight into rank
0's FC input buffer (B200_FULL_CONV_SHARD=rank0: rank 0 does all of it, the round-1 baseline).  FC6-8 run
tensor-parallel over all N GPUs through peer memory (models/alexnet_full.py: TPClassifier).
Prints V4-style lines; the "values" are the first 10 class probabilities of image 0."""
from __future__ import annotations

import os
import sys
import time

import torch

from ..config import INIT_RAND, LRN_V1, config_from_args
from ..models.alexnet_full import ConvGemm, TPClassifier, make_full_weights
from ..models.init import make_problem
from ..models.session import Session
from ..ops import ext
from ..ops.cpu import fnv1a64
from ..parallel import dist as bdist
from ..parallel.symm import DevBuf
from ..utils import logfmt as L
from .common import make_parser, fail


def batch_shard(batch: int, rank: int, world: int):
    """Images [b0, b1) of the batch whose conv3-5 + pool5 run on ``rank`` (ceil split, trailing ranks may be
    empty)."""
    per = -(-batch // world)
    b0 = min(batch, rank * per)
    return b0, min(batch, b0 + per)


def main(argv=None) -> int:
    ap = make_parser("full_alexnet", init_default=INIT_RAND, lrn_default=LRN_V1)
    ns = ap.parse_args(argv)
    cfg = config_from_args(ns, seed_default=1)
    if not torch.cuda.is_available():
        fail(bdist.env_rank_world()[0], "CUDA no CUDA-capable device is detected")
    if bdist.env_rank_world()[1] > torch.cuda.device_count():
        fail(bdist.env_rank_world()[0], "the full-network runner needs one GPU per rank")
    ctx = bdist.init()
    x, w12 = make_problem(cfg)
    sess = Session(cfg, w12, dist=ctx if ctx.world > 1 else None, device=ctx.device, precision=cfg.precision,
                   strategy=ns.strategy, scatter="sharded")
    sh = sess.engine.sh
    feat = ((sh.Hp2 - 3) // 2 + 1) * ((sh.Wp2 - 3) // 2 + 1) * 256
    w = make_full_weights(w12, seed=cfg.seed, feat=feat)
    tp = TPClassifier(w, sess.symm, ctx.rank, ctx.world, cfg.batch, feat)
    sess.symm.commit()
    shard = ctx.world > 1 and os.environ.get("B200_FULL_CONV_SHARD", "batch") == "batch"
    b0, b1 = batch_shard(cfg.batch, ctx.rank, ctx.world) if shard else (0, cfg.batch if ctx.rank == 0 else 0)
    convs = [ConvGemm(w.w3, w.b3, ctx.device, 1), ConvGemm(w.w4, w.b4, ctx.device, 1),
             ConvGemm(w.w5, w.b5, ctx.device, 1)] if b1 > b0 else None
    if shard:
        with torch.cuda.device(ctx.device):
            # rank 0's gathered LRN2 tensor and FC input, addressed through their peer mappings
            y_all = DevBuf(sess.symm.ptr(0, "out0"), cfg.batch * sh.Hp2 * sh.Wp2 * 256 * 4,
                           ctx.device).tensor((cfg.batch, sh.Hp2, sh.Wp2, 256))
            x5_all = DevBuf(sess.symm.ptr(0, "tp_x5"), cfg.batch * feat * 4, ctx.device).tensor((cfg.batch, feat))
            y_loc = torch.empty((max(b1 - b0, 1), sh.Hp2, sh.Wp2, 256), dtype=torch.float32, device=ctx.device)
    lo, hi = sess.input_rows
    xin = x[:, lo:hi].contiguous().pin_memory() if hi > lo else None

    def load():                              # this rank's input rows: pinned host -> resident device buffer
        if xin is not None:
            sess.engine.h2d(0, xin)

    def compute():
        e = sess.engine
        y = e.forward_enqueue(0, 0)
        if shard:
            tp.barrier()                     # rank 0 is past its completion wait: the gathered tensor is whole
            if b1 > b0:
                z = y_loc[:b1 - b0]
                z.copy_(y_all[b0:b1])
                for c in convs:
                    z = c(z)
                x5_all[b0:b1].copy_(ext.maxpool(z, 3, 2).reshape(b1 - b0, -1))
        elif ctx.rank == 0:
            for c in convs:
                y = c(y)
            tp.x5_view().copy_(ext.maxpool(y, 3, 2).reshape(cfg.batch, -1))
        return tp.forward()              # opens with a barrier: every rank's feature rows have landed

    def timed(fn, iters):
        torch.cuda.synchronize(ctx.device); ctx.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            out = fn()
        e1.record(); e1.synchronize()
        return out, ctx.max_float(e0.elapsed_time(e1) / iters)

    def step():
        load()
        return compute()

    ctx.barrier()
    t0 = time.perf_counter()
    probs = step()
    torch.cuda.synchronize(ctx.device)
    ctx.barrier()
    ms = (time.perf_counter() - t0) * 1e3
    kv = {}
    if ns.iters > 0:
        for _ in range(2):                   # second call captures the blocks 1-2 CUDA graph: keep it untimed
            probs = step()
        probs, dms = timed(compute, ns.iters)            # input resident on the device
        _, ems = timed(step, ns.iters)                   # plus the pinned H2D of the inputs every step
        kv = dict(device_forward_ms_warm=dms, images_per_s_warm=cfg.batch / (dms / 1e3),
                  with_h2d_forward_ms_warm=ems, with_h2d_images_per_s_warm=cfg.batch / (ems / 1e3))
    sess.engine.check_status()
    if ctx.rank == 0:
        p = probs.cpu()
        for line in L.v4_shape_values(cfg.batch, 1, 1000, p.reshape(-1)[:10].tolist()):
            L.emit(line)
        L.emit(L.v4_time(ms, "Full (spatial conv + TP FC)"))
        for line in L.extras(fnv=fnv1a64(p), wall_ms=ms, np=ctx.world, batch=cfg.batch,
                             prob_row_sum=float(p[0].sum()), argmax0=int(p[0].argmax()),
                             conv_shard="batch" if shard else "rank0", **kv):
            L.emit(line)
    ctx.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
