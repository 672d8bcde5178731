"""Per-rank device times of the real multi-GPU forward (run under torchrun / bin/mpirun, one process per GPU):
whole forward (CUDA-graph replay, ranks aligned by a device barrier before the timed loop), then this rank's block-1
and block-2 launches replayed alone (their flags are satisfied after a forward).
    HW=227 BATCH=128 ROWGROUPS=auto|rows|R HALOS=push,pull STRATEGIES=halo,recompute SCATTER=sharded"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem
from b200cnn.models.session import Session
from b200cnn.parallel import dist as bdist

ctx = bdist.init()
B = int(os.environ.get("BATCH", "128"))
HW = int(os.environ.get("HW", "227"))
RG = os.environ.get("ROWGROUPS", "auto")
rg = None if RG == "rows" else ("auto" if RG == "auto" else int(RG))
cfg = default_config(H=HW, W=HW, batch=B, init=INIT_RAND, seed=1, precision="tf32")
x, w = make_problem(cfg)
dev = ctx.device
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, n=10, do_flush=True, align=None):
    for _ in range(4): fn()
    torch.cuda.synchronize(dev); ctx.barrier()
    tot = 0.0
    for i in range(n):
        if do_flush: flush.fill_(i)
        if align: align()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    ctx.barrier()
    return tot / n * 1e3


for strategy in os.environ.get("STRATEGIES", "halo").split(","):
    for halo in os.environ.get("HALOS", "push").split(","):
        sess = Session(cfg, w, dist=ctx if ctx.world > 1 else None, device=dev, scatter=os.environ.get("SCATTER", "sharded"),
                       strategy=strategy, row_groups=rg if ctx.world > 1 else None, halo=halo)
        e = sess.engine
        xi = sess.my_input(x)
        if xi is not None:
            for i in range(2): e.input_view(i).copy_(xi.to(dev))
        torch.cuda.synchronize(dev); ctx.barrier()
        t_fwd = timeit(lambda: e.forward_enqueue(0, 0), align=e.sync_ranks)
        t_b1 = timeit(lambda: e.stage_block1(0))
        t_b2 = timeit(lambda: e.stage_block2(0))          # gate and arrival flags are already satisfied
        rows = ctx.all_gather_object((ctx.rank, round(t_fwd, 1), round(t_b1, 1), round(t_b2, 1), e.rp.n_images,
                                      e.rp.out_rows, e.rp.pool1_own))
        if ctx.is_root:
            print(f"--- world={ctx.world} ({e.plan.row_groups} row groups x {e.plan.batch_groups} batch groups) "
                  f"strategy={strategy} halo={halo} batch={B} input={HW}x{HW}  [us, L2 flushed before every launch]", flush=True)
            for r in rows:
                print(f"  rank {r[0]}: forward {r[1]:7.1f}  block1 {r[2]:7.1f}  block2 {r[3]:7.1f}  images {r[4]} out{r[5]} "
                      f"pool1{r[6]}", flush=True)
            print(f"  slowest forward {max(r[1] for r in rows):.1f} us = {B / max(r[1] for r in rows) * 1e6:,.0f} images/s; "
                  f"slowest block1+block2 {max(r[2] + r[3] for r in rows):.1f} us", flush=True)
        e.check_status()
        del sess, e
        torch.cuda.empty_cache()
ctx.finalize()
