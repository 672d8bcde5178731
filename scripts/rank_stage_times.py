"""Kernel times of every rank's share of a multi-GPU forward, measured on ONE GPU.

The single-device emulation (models/engine.py::EmulatedWorld) builds the exact launch geometry of every rank;
after one full forward all flags are satisfied, so a rank's block 1 / block 2 launch can be replayed and timed
alone.  That gives, without an N-GPU box, what each rank's kernels cost at world = 1, 2, 4, 8 — the number the
scaling curve is made of (max over ranks of block1 + block2, plus the protocol's launches) and what the tile
planner should be judged by.      WORLDS=1,2,4,8 BATCH=128 ROWGROUPS=auto python scripts/rank_stage_times.py
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem
from b200cnn.models.engine import EmulatedWorld, single_gpu_engine
from b200cnn.ops import tc

B = int(os.environ.get("BATCH", "128"))
HW = int(os.environ.get("HW", "227"))
RG = os.environ.get("ROWGROUPS", "auto")
HALO = os.environ.get("HALO", "push")
STRATEGY = os.environ.get("STRATEGY", "halo")
cfg = default_config(H=HW, W=HW, batch=B, init=INIT_RAND, seed=1, precision="tf32")
x, w = make_problem(cfg)
xd = x.cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
_a = torch.randn(4096, 4096, device="cuda")
for _ in range(60):                      # ~0.3 s of work: clocks are up before anything is timed
    _a = (_a @ _a).clamp_(-1, 1)
torch.cuda.synchronize()


def timeit(fn, n=10):
    for _ in range(4): fn()
    tot = 0.0
    for i in range(n):
        flush.fill_(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / n * 1e3


for world in [int(v) for v in os.environ.get("WORLDS", "1,2,4,8").split(",")]:
    if world == 1:
        e = single_gpu_engine(cfg, w, "cuda:0")
        e.forward(xd); torch.cuda.synchronize()
        engines = [e]
    else:
        rg = None if RG == "rows" else ("auto" if RG == "auto" else int(RG))
        ew = EmulatedWorld(cfg, w, world, "cuda:0", strategy=STRATEGY, scatter="sharded", row_groups=rg, halo=HALO)
        ew.forward(xd); torch.cuda.synchronize()
        engines = ew.engines
    plan = engines[0].plan
    print(f"== world {world}: {plan.row_groups} row groups x {plan.batch_groups} batch groups, batch {B}, "
          f"{HW}x{HW}, strategy {STRATEGY}, halo {HALO}", flush=True)
    seen, worst = set(), 0.0
    for e in engines:
        key = (e.rp.out_rows, e.rp.n_images)
        if key in seen or not e.rp.active:
            continue
        seen.add(key)
        if os.environ.get("NOSIG") == "1":          # experiment: block 1 without arrival signals / halo push
            e.rp.serves = []
        g1 = e._geom_block1(e.rank)
        g2 = e._geom_block2(e.epoch & 1)
        t1 = timeit(lambda: e.stage_block1(0))
        t2 = timeit(lambda: e.stage_block2(0))
        worst = max(worst, t1 + t2)
        print(f"  rank {e.rank}: images {e.rp.n_images:4d} out rows {e.rp.out_rows} pool1 {e.rp.pool1_own}: "
              f"block1 {t1:7.1f} us ({tc.block_tiles(g1)} tiles/img)   block2 {t2:7.1f} us "
              f"({tc.block_tiles(g2)} tiles/img)   sum {t1 + t2:7.1f} us", flush=True)
        e.check_status()
    print(f"  -> slowest rank {worst:.1f} us = {B / worst * 1e6:,.0f} images/s if nothing else cost anything", flush=True)
    del engines
    torch.cuda.empty_cache()
