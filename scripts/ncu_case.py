"""Short program for `ncu --set full` (B200_PROFILING.md: one GPU, kernel launched a few times, eager launches):
two forwards of blocks 1-2 at the bench shape (batch 128, 227x227x3, TF32) = conv1 kernel
(tc_block_swapped_kernel) + conv2 kernel (tc_block_kernel<false>), then conv3 of the full network at
batch 128 through the implicit-GEMM loader (tc_gemm_kernel<true>) twice."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem
from b200cnn.models.engine import single_gpu_engine
from b200cnn.models.alexnet_full import ConvGemm

cfg = default_config(batch=128, init=INIT_RAND, seed=1, precision="tf32")
x, w = make_problem(cfg)
e = single_gpu_engine(cfg, w, "cuda:0", "tf32")
e.use_graph = False
xd = x.cuda()
for _ in range(2):
    y = e.forward(xd)
torch.cuda.synchronize()
g = torch.Generator().manual_seed(5)
c3 = ConvGemm((torch.rand((384, 256, 3, 3), generator=g) - 0.5) * 0.04, torch.full((384,), 0.05), "cuda:0", 1,
              implicit=True)
for _ in range(2):
    y3 = c3(y)
torch.cuda.synchronize()
print("ncu_case done", tuple(y.shape), tuple(y3.shape), float(y3.abs().max()))
