"""BASELINE.json config 4 sanity on one GPU: 2048x2048x3 input (halo-dominated config), TF32 path vs the
exact-FP32 SIMT twin, plus a device-timed forward at batch 16."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem
from b200cnn.models.engine import single_gpu_engine, EmulatedWorld

cfg = default_config(H=2048, W=2048, batch=2, init=INIT_RAND, seed=1)
x, w = make_problem(cfg)
a = single_gpu_engine(cfg, w, "cuda:0", "simt").forward(x.cuda()).clone()
b = single_gpu_engine(cfg, w, "cuda:0", "tf32").forward(x.cuda()).clone()
print("shape", tuple(b.shape), "tf32 vs simt max rel", ((a - b).abs().max() / a.abs().max()).item())
ew = EmulatedWorld(cfg, w, 8, "cuda:0", precision="tf32")
c = ew.forward(x.cuda())
print("8 emulated ranks == 1 rank:", torch.equal(c, b))
del a, c, ew
cfg16 = cfg.with_(batch=16)
x16, _ = make_problem(cfg16)
e = single_gpu_engine(cfg16, w, "cuda:0", "tf32")
e.input_view(0).copy_(x16.cuda())
for _ in range(3): e.forward_enqueue(0, 0)
torch.cuda.synchronize()
t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0.record()
for _ in range(5): e.forward_enqueue(0, 0)
t1.record(); t1.synchronize()
ms = t0.elapsed_time(t1) / 5
fl = e.sh.flops_per_image() * 16
print(f"2048^2 batch 16: {ms:.3f} ms/forward, {16/ms*1e3:.1f} img/s, {fl/ms/1e9:.1f} TFLOP/s useful")
