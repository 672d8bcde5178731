"""Where conv1's time goes: block 1 re-timed with parts of the kernel switched off (Policy dbg bits of
tc_block_swapped_kernel: 64 no pixel TMA, 128 no weight TMA, 256 no epilogue work, 512 no MMAs).  The outputs are
garbage then; only the times mean something.      BATCH=128 python scripts/conv1_knobs.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem
from b200cnn.models.engine import single_gpu_engine
from b200cnn.ops import tc

B = int(os.environ.get("BATCH", "128"))
cfg = default_config(H=227, W=227, batch=B, init=INIT_RAND, seed=1, precision="tf32")
x, w = make_problem(cfg)
xd = x.cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
_a = torch.randn(4096, 4096, device="cuda")
for _ in range(60):
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


e = single_gpu_engine(cfg, w, "cuda:0")
e.forward(xd); torch.cuda.synchronize()
names = {0: "as shipped", 64: "no pixel TMA", 128: "no weight TMA", 192: "no TMA at all", 256: "no epilogue work",
         512: "no MMAs", 64 | 128 | 256: "no TMA, no epilogue (MMA issue only)", 256 | 512: "loads only",
         64 | 256 | 512: "weight loads only", 128 | 256 | 512: "pixel loads only", 4096: "no tail copies",
         4096 | 256 | 512: "TMA loads only, no tail copies", 4096 | 64 | 128 | 256: "MMA issue only, no tail copies",
         4096 | 64 | 128 | 256 | 512: "handshakes only"}
for k in [int(v) for v in os.environ.get("KNOBS", ",".join(str(k) for k in names)).split(",")]:
    tc.set_policy(dbg=k)
    t1 = timeit(lambda: e.stage_block1(0))
    print(f"dbg {k:4d}  block1 {t1:7.1f} us   {names.get(k, '')}", flush=True)
tc.set_policy(dbg=0)
e.check_status()
