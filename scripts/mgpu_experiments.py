"""Per-rank stage timings of the multi-GPU forward (run under torchrun / bin/mpirun)."""
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
cfg = default_config(H=HW, W=HW, batch=B, init=INIT_RAND, seed=1, precision="tf32")
x, w = make_problem(cfg)


def timeit(fn, n=10, sync_all=True):
    for _ in range(3): fn()
    torch.cuda.synchronize(); ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); e1.synchronize()
    ctx.barrier()
    return e0.elapsed_time(e1) / n * 1e3


for scatter in os.environ.get("SCATTERS", "p2p,sharded").split(","):
    for strategy in ("halo", "recompute"):
        sess = Session(cfg, w, dist=ctx if ctx.world > 1 else None, device=ctx.device,
                       scatter=scatter, strategy=strategy)
        e = sess.engine
        lo, hi = sess.input_rows
        if hi > lo:
            for i in range(2): e.input_view(i).copy_(x[:, lo:hi].to(ctx.device))
        torch.cuda.synchronize(); ctx.barrier()
        t_fwd = timeit(lambda: e.forward_enqueue(0, 0))
        t_b1 = timeit(lambda: e.stage_block1(0))
        t_b2 = timeit(lambda: e.stage_block2(0))         # flags already satisfied: no waiting
        def bar():
            e.stage_begin(); e.stage_barrier()
        t_bar = timeit(bar)
        rows = ctx.all_gather_object((ctx.rank, round(t_fwd, 1), round(t_b1, 1), round(t_b2, 1), round(t_bar, 1),
                                      e.rp.out_rows, e.rp.pool1_own, e.rp.in_rows))
        if ctx.is_root:
            print(f"--- world={ctx.world} scatter={scatter} strategy={strategy} batch={B} input={HW}x{HW}", flush=True)
            for r in rows:
                print(f"  rank {r[0]}: forward {r[1]} us  block1 {r[2]} us  block2 {r[3]} us  barrier {r[4]} us  "
                      f"out{r[5]} pool1{r[6]} in{r[7]}", flush=True)
        e.check_status()
        del sess, e
        torch.cuda.empty_cache()
ctx.finalize()
