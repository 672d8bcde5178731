"""Full AlexNet forward (extension; BASELINE.json config 5), one process per GPU.

    mpirun -np N python -m b200cnn.runners.full --batch 256

Blocks 1-2 run decomposed across the N GPUs (rows, or ``--row-groups``: rows x batch) with the fused peer-memory
kernels, every rank KEEPING the rows it computed (gather="local").  Conv3-5 + Pool5 work on 13x13 maps, too small to
split by rows, so they are sharded by BATCH inside each group of ranks that split one image set: after a device-side
barrier a rank collects, for its share of the images, the row blocks the other members computed (one pass of peer
reads over NVLink), runs the three implicit-GEMM convolutions and the pool locally and keeps its feature rows.
Nothing is funnelled through rank 0: FC6 reads every rank's feature rows from where they are (all-gather fused into
the GEMM loader), FC7's partial products are added by the shard owners in rank order (deterministic reduce-scatter
by pull), FC8 writes its logits columns into rank 0's buffer (models/alexnet_full.py: TPClassifier).
Prints V4-style lines; the "values" are the first 10 class probabilities of image 0."""
from __future__ import annotations

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
from .common import make_parser, fail, weight_flags


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
    loaded = None
    if ns.load_weights:                      # a full-network file: every rank maps it and reads only what it uses
        from ..utils import checkpoint as ck
        try:
            loaded = ck.load_weights(ns.load_weights, mmap=True)
            ck.check_against(cfg, loaded)
            if not hasattr(loaded, "blocks12"):
                raise ValueError("a blocks-1-2 file; the full network needs kind 'full'")
        except (OSError, ValueError) as ex:
            fail(ctx.rank, f"--load-weights {ns.load_weights}: {ex}")
        w12 = loaded.blocks12
    from .v4 import _row_groups
    from ..parallel.planner import split_rows
    sess = Session(cfg, w12, dist=ctx if ctx.world > 1 else None, device=ctx.device, precision=cfg.precision,
                   strategy=ns.strategy, scatter="sharded", row_groups=_row_groups(ns), gather="local")
    e = sess.engine
    sh, plan, rp = e.sh, e.plan, e.rp
    feat = ((sh.Hp2 - 3) // 2 + 1) * ((sh.Wp2 - 3) // 2 + 1) * 256
    if loaded is not None:
        if tuple(loaded.w6.shape) != (4096, feat):
            fail(ctx.rank, f"--load-weights {ns.load_weights}: w6 has shape {tuple(loaded.w6.shape)}, "
                           f"this input size needs (4096, {feat})")
        w = loaded
    else:
        w = make_full_weights(w12, seed=cfg.seed, feat=feat)
    if ns.save_weights and ctx.rank == 0:
        from ..utils import checkpoint as ck
        ck.save_weights(ns.save_weights, w, meta={"seed": int(cfg.seed), "init": str(cfg.init), "feat": int(feat)})
    tp = TPClassifier(w, sess.symm, ctx.rank, ctx.world, cfg.batch, feat)
    sess.symm.commit()
    # Conv3-5 + pool5 by batch: the ranks that split the rows of one image set share its images out among themselves;
    # a rank collects, for ITS images, the row blocks the other members computed (peer reads of their output slabs)
    def share_of(q):
        members = [m for m in plan.ranks if m.batch_group == q.batch_group]
        lo, hi = split_rows(q.n_images, len(members))[q.row_group]
        return q.images[0] + lo, q.images[0] + hi
    b0, b1 = share_of(rp)
    tp.x5_sources = [(q.rank,) + share_of(q) for q in plan.ranks]
    members = [m for m in plan.ranks if m.batch_group == rp.batch_group and m.out_rows[1] > m.out_rows[0]]
    convs = [ConvGemm(w.w3, w.b3, ctx.device, 1), ConvGemm(w.w4, w.b4, ctx.device, 1),
             ConvGemm(w.w5, w.b5, ctx.device, 1)] if b1 > b0 else None
    row_bytes = sh.Wp2 * 256 * 4
    with torch.cuda.device(ctx.device):
        z_in = torch.empty((max(b1 - b0, 1), sh.Hp2, sh.Wp2, 256), dtype=torch.float32, device=ctx.device)
        parts = []                           # (rows, view of the member's output slab restricted to my images)
        for m in members:
            r0, r1 = m.out_rows
            if ctx.world == 1:
                break
            slab = DevBuf(sess.symm.ptr(m.rank, "out0"), m.n_images * (r1 - r0) * row_bytes,
                          ctx.device).tensor((m.n_images, r1 - r0, sh.Wp2, 256))
            parts.append((r0, r1, slab[b0 - m.images[0]:b1 - m.images[0]]))
        x5_mine = tp.x5_view()[b0:b1]
    xin = sess.my_input(x)
    xin = xin.contiguous().pin_memory() if xin is not None else None

    def load():                              # this rank's input rows: pinned host -> resident device buffer
        if xin is not None:
            e.h2d(0, xin)

    def compute():
        y = e.forward_enqueue(0, 0)
        if ctx.world > 1:
            tp.barrier()                     # every member's block 2 is finished: its output slab is whole
        if b1 > b0:
            if ctx.world > 1:
                for r0, r1, src in parts:
                    z_in[:b1 - b0, r0:r1].copy_(src)
                z = z_in[:b1 - b0]
            else:
                z = y
            for c in convs:
                z = c(z)
            x5_mine.copy_(ext.maxpool(z, 3, 2).reshape(b1 - b0, -1))
        return tp.forward()                  # opens with a barrier: every rank's feature rows are in place

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
    e.check_status()
    if ctx.rank == 0:
        p = probs.cpu()
        for line in L.v4_shape_values(cfg.batch, 1, 1000, p.reshape(-1)[:10].tolist()):
            L.emit(line)
        L.emit(L.v4_time(ms, "Full (spatial conv + TP FC)"))
        for line in L.extras(fnv=fnv1a64(p), wall_ms=ms, np=ctx.world, batch=cfg.batch,
                             prob_row_sum=float(p[0].sum()), argmax0=int(p[0].argmax()),
                             decomposition=f"{plan.row_groups}x{plan.batch_groups}", **kv):
            L.emit(line)
    ctx.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
