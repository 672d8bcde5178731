#!/usr/bin/env python
"""Per-stage CUDA-event times of the full-AlexNet extension on one GPU (config 5 shape: batch 256), with
conv3-5 through the materialised im2col and through the implicit-GEMM loader.
Usage: python scripts/full_stage_times.py [batch] [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from b200cnn import default_config
from b200cnn.config import INIT_RAND
from b200cnn.models.alexnet_full import ConvGemm, FullAlexNet, make_full_weights
from b200cnn.models.init import make_problem
from b200cnn.ops import ext


def timed(fn, reps):
    fn(); torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in ev)
    return t[len(t) // 2] * 1e3


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 7
    cfg = default_config(batch=batch, init=INIT_RAND, seed=2, precision="tf32")
    x, w12 = make_problem(cfg)
    w = make_full_weights(w12, seed=5)
    net = FullAlexNet(cfg, w, "cuda:0", tp_world=1)
    xd = x.cuda()
    y2 = net.engine.forward(xd).clone()
    print(f"batch {batch}: blocks 1-2 {timed(lambda: net.engine.forward(xd), reps):9.1f} us")
    outs = {}
    for implicit in (False, True):
        c3, c4, c5 = (ConvGemm(a, b, "cuda:0", 1, implicit=implicit) for a, b in
                      ((w.w3, w.b3), (w.w4, w.b4), (w.w5, w.b5)))
        y3 = c3(y2); y4 = c4(y3); y5 = c5(y4)
        outs[implicit] = y5
        tag = "implicit" if implicit else "im2col  "
        t3, t4, t5 = timed(lambda: c3(y2), reps), timed(lambda: c4(y3), reps), timed(lambda: c5(y4), reps)
        gf = lambda ci, co: 2.0 * batch * 169 * 9 * ci * co / 1e9
        print(f"conv3-5 {tag}: {t3:8.1f} {t4:8.1f} {t5:8.1f} us = "
              f"{gf(256, 384) / t3 * 1e3:6.0f} {gf(384, 384) / t4 * 1e3:6.0f} {gf(384, 256) / t5 * 1e3:6.0f} TFLOP/s")
        if not implicit:
            print(f"   of which im2col alone: {timed(lambda: ext.im2col(y2, 3, 1, 1), reps):8.1f} "
                  f"{timed(lambda: ext.im2col(y3, 3, 1, 1), reps):8.1f} us")
    print("implicit == im2col:", torch.equal(outs[False], outs[True]))
    y5 = outs[True]
    print(f"pool5 {timed(lambda: ext.maxpool(y5, 3, 2), reps):8.1f} us")
    feat = ext.maxpool(y5, 3, 2).reshape(batch, -1)
    t = net.tp[0]
    t.x5_view().copy_(feat)

    def fc():
        t.phase_fc6(); t.phase_fc7(); t.phase_fc7_finish(); t.phase_fc8(); t.phase_softmax()
    print(f"fc6-8 + softmax {timed(fc, reps):8.1f} us")
    print(f"whole forward (default path) {timed(lambda: net.forward(xd), reps):9.1f} us")


if __name__ == "__main__":
    main()
