"""Small case for compute-sanitizer (one tool per gpurun call, B200_PROFILING.md):
    compute-sanitizer --tool memcheck python scripts/sanitize_case.py
Covers the SIMT and tcgen05 block kernels (both conv1 loader modes), the flag kernels, the 2-rank
emulation (peer segments + signals), the NCHW adaptor (both entry points) and the GEMM/extension kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem
from b200cnn.models.engine import single_gpu_engine, EmulatedWorld
from b200cnn.ops import cuda as cu, ext

cfg = default_config(H=99, W=99, batch=1, init=INIT_RAND, seed=1)
x, w = make_problem(cfg)
for prec in ("simt", "tf32"):
    y = single_gpu_engine(cfg, w, "cuda:0", prec).forward(x.cuda())
    torch.cuda.synchronize()
    print(prec, tuple(y.shape), float(y.abs().max()))
y2 = EmulatedWorld(cfg, w, 2, "cuda:0", precision="tf32").forward(x.cuda())
print("emulated np2 equal:", torch.equal(y2, y))
xn = torch.randn(1, 3, 31, 31).cuda(); wn = torch.randn(8, 3, 3, 3).cuda()
print("conv_nchw", tuple(cu.conv_nchw(xn, wn, None, 1, 1).shape))
a = torch.randn(70, 64).cuda(); lin = ext.TcLinear(torch.randn(130, 64), "cuda:0")
print("gemm", tuple(lin(a, torch.zeros(130).cuda(), relu=True).shape))
print("softmax", float(ext.softmax(torch.randn(3, 10).cuda()).sum()))
torch.cuda.synchronize()
print("sanitize_case done")
