"""Per-stage device times of one forward on one GPU, warm-L2 and with a 256 MiB L2 flush before each stage."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem
from b200cnn.models.engine import single_gpu_engine

B = int(os.environ.get("BATCH", "128"))
cfg = default_config(batch=B, init=INIT_RAND, seed=1, precision="tf32")
x, w = make_problem(cfg)
eng = single_gpu_engine(cfg, w, "cuda:0")
eng.input_view(0).copy_(x.cuda())
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

def timeit(fn, do_flush, n=10):
    for _ in range(3): fn()
    tot = 0.0
    for i in range(n):
        if do_flush: flush.fill_(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / n * 1e3

stages = [("s2d", lambda: eng.stage_s2d(0)), ("block1", lambda: eng.stage_block1(0)), ("block2", lambda: eng.stage_block2(0)),
          ("forward(eager stages)", lambda: (eng.stage_s2d(0), eng.stage_block1(0), eng.stage_block2(0)))]
for name, fn in stages:
    print(f"{name:24s} warm {timeit(fn, False):8.1f} us   flushed {timeit(fn, True):8.1f} us", flush=True)
