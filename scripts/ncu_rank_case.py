"""Short program for `ncu --set full`: the block-1 and block-2 launches of ONE rank of an emulated N-GPU forward
(batch 128, 227x227x3), replayed after a full forward so every flag is satisfied.  WORLD=8 (default) is the
launch-bound case: 32 images x half the rows per rank.  WORLD=1 profiles the single-GPU kernels.
    WORLD=8 RANK=0 python scripts/ncu_rank_case.py
    ncu --set full --clock-control none --import-source on -k regex:'tc_block' -s <skip> -c 2 -o gpurun_out/prof ...
The script prints how many tc_block launches precede the replayed ones (the value for ncu's -s)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem
from b200cnn.models.engine import EmulatedWorld, single_gpu_engine

W = int(os.environ.get("WORLD", "8"))
R = int(os.environ.get("RANK", "0"))
cfg = default_config(batch=128, init=INIT_RAND, seed=1, precision="tf32")
x, w = make_problem(cfg)
xd = x.cuda()
if W == 1:
    e = single_gpu_engine(cfg, w, "cuda:0"); e.use_graph = False
    e.forward(xd); torch.cuda.synchronize()
    skip = 2
else:
    ew = EmulatedWorld(cfg, w, W, "cuda:0", strategy="halo", scatter="sharded", row_groups="auto", halo="push")
    ew.forward(xd); torch.cuda.synchronize()
    e = ew.engines[R]
    skip = 2 * sum(1 for q in ew.engines if q.rp.active)
for _ in range(2):
    e.stage_block1(0)
    e.stage_block2(0)
torch.cuda.synchronize()
e.check_status()
print(f"ncu_rank_case world {W} rank {R}: skip {skip} tc_block launches, then [block1, block2] x 2")
