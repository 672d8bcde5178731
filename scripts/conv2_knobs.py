"""Where conv2's time goes: block 2 (channel-split pair kernel) re-timed with parts switched off (Policy dbg bits:
64 no pixel TMA, 128 no weight TMA, 256 no epilogue work, 512 no MMAs; 2048 at plan time = pool through the smem
strips instead of registers).  Outputs are garbage then; only the times mean something."""
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
names = {0: "as shipped", 2048: "pooling through smem strips", 64: "no pixel TMA", 128: "no weight TMA",
         192: "no TMA", 256: "no epilogue work", 512: "no MMAs", 64 | 128 | 256: "MMA issue only",
         256 | 512: "loads only", 64 | 128 | 256 | 512: "handshakes only"}
for k in [int(v) for v in os.environ.get("KNOBS", ",".join(str(k) for k in names)).split(",")]:
    tc.set_policy(dbg=k)
    t2 = timeit(lambda: e.stage_block2(0))
    print(f"dbg {k:4d}  block2 {t2:7.1f} us   {names.get(k, '')}", flush=True)
tc.set_policy(dbg=0)
e.check_status()
