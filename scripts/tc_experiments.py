"""Perf decomposition of tc_block on a GPU box: time block1/block2 with parts of the pipeline
disabled (B200_TC_DBG bits: 1 no A copies, 2 no B TMA, 4 no epilogue math).  Results are garbage
numerically; only the timings matter."""
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
eng.stage_s2d(0)

def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); e1.synchronize()
    return e0.elapsed_time(e1) / n * 1e3

for dbg in (0,):
    os.environ["B200_TC_DBG"] = str(dbg)
    eng.epoch = 0
    t1 = timeit(lambda: eng.stage_block1(0))
    t2 = timeit(lambda: eng.stage_block2(0))
    print(f"dbg={dbg} (A{'-' if dbg&1 else '+'} B{'-' if dbg&2 else '+'} epi{'-' if dbg&4 else '+'} {'MMA-free-run' if dbg&8 else ''}{' noLRN' if dbg&16 else ''}{' noPool' if dbg&32 else ''}{' noLRNstore' if dbg&64 else ''}{' noLRNmath' if dbg&128 else ''}): "
          f"block1 {t1:8.1f} us   block2 {t2:8.1f} us", flush=True)
