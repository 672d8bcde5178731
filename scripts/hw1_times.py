"""HW1 equivalent (FP64 row-block GEMM, csrc/cuda/dgemm.cu) on one GPU: device time and GFLOP/s of b200_dgemm_rows at
several n, against cuBLAS DGEMM on the same box (torch.matmul, float64) as the measured FP64 ceiling — the numbers
behind the K10 row of profiles/ROOFLINE.md.  Multi-GPU HW1 times come from the runner itself:
    bin/mpirun -np N python -m b200cnn.runners.hw1 4096 --backend gpu
        python scripts/hw1_times.py [n ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200cnn.ops import cuda as cu


def timeit(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); e1.synchronize()
    return e0.elapsed_time(e1) / n


for n in [int(a) for a in sys.argv[1:]] or [1024, 2048, 4096]:
    g = torch.Generator().manual_seed(n)
    A = torch.randint(0, 10, (n, n), generator=g).double().cuda()
    B = torch.randint(0, 10, (n, n), generator=g).double().cuda()
    C = torch.empty((n, n), dtype=torch.float64, device="cuda")
    ms = timeit(lambda: cu.dgemm_rows(A.data_ptr(), B.data_ptr(), C.data_ptr(), n, n))
    ok = torch.equal(C, A @ B)
    ms_ref = timeit(lambda: torch.matmul(A, B))
    flop = 2.0 * n ** 3
    print(f"n={n}: b200_dgemm_rows {ms:.3f} ms = {flop / ms / 1e6:,.0f} GFLOP/s (exact: {ok});  cuBLAS DGEMM {ms_ref:.3f} ms = "
          f"{flop / ms_ref / 1e6:,.0f} GFLOP/s;  ratio {ms_ref / ms:.2f}", flush=True)
