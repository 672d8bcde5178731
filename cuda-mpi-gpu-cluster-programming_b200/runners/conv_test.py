"""conv_test — stand-alone convolution check (ref final_project/conv_test: a prebuilt binary whose
sources are missing; recovered behaviour: NCHW direct convolution, prints
``Convolution Test Output (first 10 values):``).

Runs the framework's NCHW adaptor (csrc/cuda/conv_nchw.cu) on a deterministic ramp input and, since
the reference program checks nothing, additionally compares against the CPU oracle and prints
PASSED/FAILED after the reference's line."""
from __future__ import annotations

import argparse
import sys

import torch

from ..ops import cpu as cpu_ops
from ..utils import logfmt as L


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="conv_test")
    for name, dv in (("n", 1), ("c", 3), ("h", 32), ("w", 32), ("k", 8), ("f", 3), ("s", 1), ("p", 0)):
        ap.add_argument(f"--{name}", type=int, default=dv)
    ap.add_argument("--cpu", action="store_true", help="oracle only (no GPU)")
    ns = ap.parse_args(argv)
    x = (torch.arange(ns.n * ns.c * ns.h * ns.w, dtype=torch.float32).reshape(ns.n, ns.c, ns.h, ns.w) % 17) * 0.1
    w = ((torch.arange(ns.k * ns.c * ns.f * ns.f, dtype=torch.float32).reshape(ns.k, ns.c, ns.f, ns.f) % 7) - 3) * 0.01
    b = torch.zeros(ns.k)
    ref, _ = cpu_ops.conv(x.permute(0, 2, 3, 1).contiguous(), w, b, ns.f, ns.s, ns.p)
    ref = ref.permute(0, 3, 1, 2).contiguous()
    if ns.cpu or not torch.cuda.is_available():
        out, how = ref, "cpu-oracle"
    else:
        from ..ops import cuda as cu
        out, how = cu.conv_nchw(x.cuda(), w.cuda(), b.cuda(), ns.s, ns.p).cpu(), "gpu"
    flat = out.reshape(-1)
    L.emit("Convolution Test Output (first 10 values): " + " ".join(L.cxx_float(v) for v in flat[:10].tolist()))
    err = (out - ref).abs().max().item() / max(ref.abs().max().item(), 1e-30)
    L.emit(f"Output Shape (NCHW): {'x'.join(str(s) for s in out.shape)}")
    L.emit(f"max_rel_err_vs_cpu_oracle={L.cxx_float(err)} backend={how}")
    L.emit("Convolution Test: " + ("PASSED" if err < 1e-5 else "FAILED"))
    return 0 if err < 1e-5 else 1


if __name__ == "__main__":
    sys.exit(main())
