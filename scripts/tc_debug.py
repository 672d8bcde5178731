"""Per-block diagnostics of the tcgen05 path against the fp32 torch reference (run on a GPU box).
Prints where errors sit (by channel / row / column) so one gpurun call localises a layout bug."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.init import make_problem
from b200cnn.models.alexnet import forward_torch_reference
from b200cnn.models.engine import single_gpu_engine


def report(name, got, ref):
    d = (got - ref).abs()
    scale = ref.abs().max().item()
    print(f"[{name}] shape {tuple(got.shape)} max|ref| {scale:.4g} max err {d.max().item():.4g} "
          f"rel {d.max().item()/scale:.3e} mean err {d.mean().item():.3e} nan {torch.isnan(got).sum().item()}")
    if d.max().item() / scale > 4e-3:
        bad = d > 4e-3 * scale
        print("   bad frac", bad.float().mean().item())
        print("   bad by channel (first 40):", bad.float().mean(dim=(0, 1, 2))[:40].tolist())
        print("   bad by row:", bad.float().mean(dim=(0, 2, 3)).tolist())
        print("   bad by col:", bad.float().mean(dim=(0, 1, 3)).tolist())
        print("   bad by image:", bad.float().mean(dim=(1, 2, 3)).tolist())
        idx = bad.nonzero()[:8]
        for i in idx:
            i = tuple(i.tolist())
            print("   e.g.", i, "got", got[i].item(), "ref", ref[i].item())


for (H, W, B) in [(227, 227, 2), (99, 131, 1)]:
    cfg = default_config(H=H, W=W, batch=B, init=INIT_RAND, seed=5, precision="tf32")
    x, w = make_problem(cfg)
    eng = single_gpu_engine(cfg, w, "cuda:0")
    y = eng.forward(x.cuda())
    torch.cuda.synchronize()
    print("status", eng.status())
    p1 = eng.pool1_view().cpu()
    report(f"pool1 {H}x{W}", p1, forward_torch_reference(x, w, cfg, upto="pool1"))
    report(f"out   {H}x{W}", y.cpu(), forward_torch_reference(x, w, cfg))
print("tc_debug done")
