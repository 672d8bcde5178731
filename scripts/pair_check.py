"""CTA-pair conv2 kernel (B200_TC_PAIR=1, draft: cta_group::2, TMA-fed operands) vs the default kernel: same output bits, block times."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem
from b200cnn.models.engine import single_gpu_engine


def run(batch, db, reps):  # db: use the pair kernel
    if db:
        os.environ["B200_TC_PAIR"] = "1"
    else:
        os.environ.pop("B200_TC_PAIR", None)
    cfg = default_config(batch=batch, init=INIT_RAND, seed=1, precision="tf32")
    x, w = make_problem(cfg)
    e = single_gpu_engine(cfg, w, "cuda:0", "tf32")
    e.use_graph = False
    y = e.forward(x.cuda()).clone()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); e.forward_enqueue(0, 0); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return y, sorted(ts)[len(ts) // 2] if ts else 0.0


y0, _ = run(6, False, 0)
y1, _ = run(6, True, 0)
print("batch 6: equal", torch.equal(y0, y1), "maxdiff", float((y0 - y1).abs().max()), flush=True)
_, t0 = run(128, False, 5)
_, t1 = run(128, True, 5)
print(f"batch 128 forward (conv1+conv2, L2 warm): default {t0:.1f} us, CTA-pair conv2 {t1:.1f} us", flush=True)
