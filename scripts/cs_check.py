"""Channel-split CTA-pair conv2 kernel (policy conv2="cs") against the double-buffered kernel: output bits on several
shapes, then block-2 times at batch 128 for world 1 and for the ranks of an emulated 8-GPU run.  Own process on
purpose (a new tcgen05/TMA form that faults must not take a pytest session with it):
    timeout 120 python scripts/cs_check.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem
from b200cnn.models.engine import single_gpu_engine, EmulatedWorld
from b200cnn.ops import tc


def fwd(cfg, x, w, kind, reuse=0):
    tc.set_policy(conv2=kind)
    e = single_gpu_engine(cfg, w, "cuda:0", "tf32")
    e.use_graph = False
    y = e.forward(x.cuda()).clone()
    y2 = e.forward(x.cuda()).clone()
    e.check_status()
    return y, y2


for batch, hw in ((6, 227), (3, 99)):
    cfg = default_config(H=hw, W=hw, batch=batch, init=INIT_RAND, seed=2, precision="tf32")
    x, w = make_problem(cfg)
    a, _ = fwd(cfg, x, w, "db")
    b, b2 = fwd(cfg, x, w, "cs")
    rel = float((a - b).abs().max() / a.abs().max())
    print(f"batch {batch} {hw}x{hw}: equal {torch.equal(a, b)} rel diff {rel:.3e} second forward equal {torch.equal(b, b2)}",
          flush=True)

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn, n=10):
    for _ in range(4): fn()
    tot = 0.0
    for i in range(n):
        flush.fill_(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / n * 1e3


cfg = default_config(batch=128, init=INIT_RAND, seed=1, precision="tf32")
x, w = make_problem(cfg)
xd = x.cuda()
for kind in ("pair", "cs"):
    tc.set_policy(conv2=kind)
    e = single_gpu_engine(cfg, w, "cuda:0")
    e.forward(xd); torch.cuda.synchronize()
    print(f"world 1 conv2={kind}: block2 {timeit(lambda: e.stage_block2(0)):.1f} us", flush=True)
    for world in (2, 4, 8):
        ew = EmulatedWorld(cfg, w, world, "cuda:0", strategy="halo", scatter="sharded", row_groups="auto", halo="push")
        y = ew.forward(xd); torch.cuda.synchronize()
        ts = [timeit(lambda e=e: e.stage_block2(0)) for e in ew.engines[:2]]
        print(f"world {world} conv2={kind}: block2 rank0 {ts[0]:.1f} us rank1 {ts[1]:.1f} us", flush=True)
        del ew
tc.set_policy(conv2="auto")
