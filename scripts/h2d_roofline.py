"""What the host links allow for the end-to-end number: every rank copies its share of one step's input
(pinned host -> device) and of one step's result (device -> pinned host) at the same time as all other ranks,
with nothing else running: what the host links give a plain copy loop (one stream per direction and rank, copies back
to back).  bench.py's `e2e` is bound by the same links; with three steps in flight it lands within a few percent of this
number and at 8 GPUs slightly above it (the files under profiles/r2 call the field `e2e_ceiling_images_per_s`).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29720 \
        scripts/h2d_roofline.py            (N = 1: plain `python scripts/h2d_roofline.py`)

Bytes per step are bench.py's (batch 128, 227x227x3 in, 13x13x256 out): each rank gets 1/N of both.
"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

IN_BYTES = 128 * 227 * 227 * 3 * 4
OUT_BYTES = 128 * 13 * 13 * 256 * 4


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n_in, n_out = IN_BYTES // world // 4, OUT_BYTES // world // 4
    h_in = torch.empty(n_in, dtype=torch.float32).pin_memory()
    h_out = torch.empty(n_out, dtype=torch.float32).pin_memory()
    d_in = torch.empty(n_in, dtype=torch.float32, device="cuda")
    d_out = torch.zeros(n_out, dtype=torch.float32, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    reps = int(os.environ.get("REPS", "40"))
    res = {}
    for mode in ("h2d", "d2h", "both"):
        for timed in (False, True):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(reps if timed else 5):
                if mode in ("h2d", "both"):
                    with torch.cuda.stream(s1):
                        d_in.copy_(h_in, non_blocking=True)
                if mode in ("d2h", "both"):
                    with torch.cuda.stream(s2):
                        h_out.copy_(d_out, non_blocking=True)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res[mode] = float(t.item()) / reps
    if rank == 0:
        h2d = IN_BYTES / res["h2d"] / 1e9
        d2h = OUT_BYTES / res["d2h"] / 1e9
        both = (IN_BYTES + OUT_BYTES) / res["both"] / 1e9
        print(json.dumps({
            "n_gpus": world, "h2d_bytes_per_step": IN_BYTES, "d2h_bytes_per_step": OUT_BYTES,
            "aggregate_h2d_GBps": round(h2d, 1), "aggregate_d2h_GBps": round(d2h, 1),
            "aggregate_both_directions_GBps": round(both, 1),
            "step_ms_copies_only": round(res["both"] * 1e3, 4),
            "copy_loop_images_per_s": round(128 / res["both"], 1),
            "note": "slowest rank, wall clock over %d back-to-back steps of copies, nothing else running" % reps}),
            flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    sys.exit(main())
