#!/usr/bin/env python
"""Headline benchmark: AlexNet blocks 1-2 forward, images/s, row-decomposed over N B200s.

Config (BASELINE.json #3): 227x227x3 fp32 inputs, batch 128, synthetic data (V1 rand() stream),
random-init weights, spatial row decomposition with peer-memory scatter/halo/gather, TF32
tensor-core kernels (the reference computes in fp32; --precision simt gives the exact-fp32 twin).

    python bench.py [--gpus N --steps K --warmup W]              # our framework
    python bench.py --impl reference [...]                       # unmodified reference (baseline/_ref)

For N>1 launch with torchrun (one rank per GPU).  Rank 0 prints ONE JSON line.
  value      device-timed images/s: CUDA events around each forward on the launching stream, L2
             flushed between timed iterations (outside the events), max over ranks.
  e2e.value  images/s through the public API (b200cnn.models.session.Session): every step copies
             its input from pinned host memory, runs the forward, and copies the result back.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

sys.path.insert(0, os.path.join(ROOT, "baseline"))
from bench_contract import METRIC, UNIT, HIGHER_IS_BETTER, SCALING, MODEL  # noqa: E402  (shared with the reference arm)

# Reference's published best V4 images/s per np (best_runs.csv:10-12; other HW).
REF_V4_BEST_IMG_S = {1: 1 / 0.183, 2: 1 / 0.181, 4: 1 / 0.250}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--height", type=int, default=227)
    ap.add_argument("--width", type=int, default=227)
    ap.add_argument("--precision", choices=["tf32", "simt"], default="tf32")
    ap.add_argument("--strategy", choices=["halo", "recompute"], default="halo",
                    help="halo (default): per-layer row ownership, boundary pool1 rows cross NVLink inside the fused "
                         "kernels; recompute: overlapped tiling, no inter-layer traffic")
    ap.add_argument("--halo", choices=["push", "pull"], default="push",
                    help="push: conv1's epilogue stores boundary pool1 rows into the neighbour's slab (P2P stores), conv2 "
                         "reads local memory through TMA; pull: conv2's loader reads them from the neighbour (P2P loads)")
    ap.add_argument("--row-groups", default="auto",
                    help="auto: 2 row groups x N/2 batch groups; rows: pure row split over all N ranks; or a divisor of N")
    ap.add_argument("--scatter", choices=["p2p", "sharded"], default="sharded",
                    help="where the images live.  sharded = every rank holds (and, end to end, uploads) the rows of the "
                         "images it works on; p2p = all images on rank 0, pulled over NVLink inside conv1")
    ap.add_argument("--e2e-gather", choices=["host", "rank0"], default="host",
                    help="end-to-end run: host = every rank copies its rows into one shared pinned host tensor; rank0 = "
                         "rows are stored into rank 0's device tensor over NVLink and rank 0 copies all of it")
    ap.add_argument("--alt", action="store_true",
                    help="also time the alternatives (p2p scatter, pure rows, recompute) for config.alternatives")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-check", action="store_true", help="skip the comparison with a one-rank forward")
    return ap.parse_args()


def _row_groups(a, N):
    if N == 1 or a.row_groups == "rows":
        return None
    return "auto" if a.row_groups == "auto" else int(a.row_groups)


def run_ours(a):
    import torch
    import b200cnn
    from b200cnn.config import INIT_RAND
    from b200cnn.models.init import make_problem
    from b200cnn.models.session import Session
    from b200cnn.parallel import dist as bdist
    from b200cnn.utils.clocks import ClockSampler, rejected

    if not torch.cuda.is_available():       # never a silent CPU fallback: the metric is a GPU metric
        print(json.dumps({"impl": "ours", "error": "no CUDA device visible: bench.py measures the sm_100a kernels "
                                                   "and has no CPU path", "gpu_launches": 0}))
        sys.exit(2)
    ctx = bdist.init()
    if ctx.world != a.gpus and ctx.is_root:
        sys.stderr.write(f"bench: --gpus {a.gpus} but WORLD_SIZE={ctx.world}; using WORLD_SIZE\n")
    N = ctx.world
    precision = a.precision
    if precision == "tf32":
        from b200cnn.ops import tc  # noqa: F401   (raises if the tcgen05 kernels are not in the library)
    cfg = b200cnn.default_config(H=a.height, W=a.width, batch=a.batch, init=INIT_RAND, seed=1,
                                 precision=precision)
    x, w = make_problem(cfg)             # synthetic inputs, random-init weights (same on every rank)
    dev = ctx.device
    torch.cuda.set_device(dev)
    rg = _row_groups(a, N)
    dist_arg = ctx if N > 1 else None

    def load_inputs(sess, nbuf):
        xi = sess.my_input(x)
        if xi is not None:
            for i in range(nbuf):
                sess.engine.input_view(i).copy_(xi.to(dev))

    def check_against_one_rank(sess):
        """Rank 0's gathered output must be bit-equal to a one-rank forward of the same input (the property
        the reference's V2.2/V4 decomposition does not have).  Returns (ok, fnv)."""
        from b200cnn.models.engine import single_gpu_engine
        from b200cnn.ops.cpu import fnv1a64
        load_inputs(sess, 1)
        torch.cuda.synchronize(dev)
        ctx.barrier()
        y = sess.engine.forward_enqueue(0, 0)
        sess.engine.check_status()
        ok, fnv = True, None
        if ctx.is_root:
            y = y.clone()
            one = single_gpu_engine(cfg, w, dev, precision)
            ref = one.forward(x.to(dev))
            one.check_status()
            ok = bool(torch.equal(ref, y))
            fnv = f"0x{fnv1a64(y.cpu()):016x}"
            del one, ref
            torch.cuda.empty_cache()
        ctx.barrier()
        return ctx.broadcast_object((ok, fnv), 0)

    def measure_device(sess):
        e = sess.engine
        load_inputs(sess, 2)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2
        for i in range(max(a.warmup, 6)):          # >= 2 calls per buffer parity: both CUDA graphs are
            e.forward_enqueue(i % 2, i % 2)        # captured here, never inside the timed region
        torch.cuda.synchronize(dev)
        ctx.barrier()
        torch.cuda.synchronize(dev)
        sampler = ClockSampler(dev.index or 0).start()
        l0 = e.launches
        evs = []
        for i in range(a.steps):                 # enqueue everything, sync once: launch overhead of
            flush.fill_(i & 0xFF)                # the next step hides behind the (untimed) L2 flush
            e.sync_ranks()                       # untimed: every rank has finished its flush
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            e.forward_enqueue(i % 2, i % 2)
            ev1.record()
            evs.append((ev0, ev1))
        torch.cuda.synchronize(dev)
        total_ms = sum(a_.elapsed_time(b_) for a_, b_ in evs)
        ctx.barrier()
        clocks = sampler.stop()
        launches = e.launches - l0
        e.check_status()
        del flush
        return ctx.max_float(total_ms), launches, clocks

    def quick_ms(**kw):
        """Warm device time per forward of an alternative configuration (no L2 flush; for config.alternatives)."""
        s_ = Session(cfg, w, dist=dist_arg, device=dev, precision=precision, **kw)
        load_inputs(s_, 1)
        e_ = s_.engine
        for _ in range(4):
            e_.forward_enqueue(0, 0)
        torch.cuda.synchronize(dev)
        ctx.barrier()
        e_.sync_ranks()
        t0_, t1_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0_.record()
        for _ in range(6):
            e_.forward_enqueue(0, 0)
        t1_.record()
        t1_.synchronize()
        ms_ = ctx.max_float(t0_.elapsed_time(t1_) / 6)
        e_.check_status()
        del s_, e_
        torch.cuda.empty_cache()
        return round(a.batch / (ms_ / 1e3), 1)

    def measure_e2e(sess):
        """Wall clock over K steps through Session.submit/result: every step uploads its input from pinned host
        memory and brings its result back to (pinned) host memory; <= 3 steps in flight, results consumed in order."""
        xi = sess.my_input(x)
        hx = [xi.contiguous().pin_memory() for _ in range(2)] if xi is not None else None
        for i in range(max(a.warmup, 2 * sess.slots)):
            sess.result(sess.submit(hx[i % 2] if hx else None))
        sess.drain()
        ctx.barrier()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        tickets, last = [], None
        for i in range(a.steps):
            tickets.append(sess.submit(hx[i % 2] if hx else None))
            if len(tickets) >= sess.slots - 1:
                last = sess.result(tickets.pop(0))
        for t in tickets:
            last = sess.result(t)
        sess.drain()
        torch.cuda.synchronize(dev)
        ms = (time.perf_counter() - t0) * 1e3
        sess.engine.check_status()
        return ctx.max_float(ms), sess.h2d_bytes, sess.d2h_bytes, last

    strategy = a.strategy if N > 1 else "single-gpu"
    scatter = a.scatter if N > 1 else "local"
    sess = Session(cfg, w, dist=dist_arg, device=dev, precision=precision, strategy=a.strategy,
                   scatter=a.scatter, row_groups=rg, gather="rank0", halo=a.halo)
    plan = sess.engine.plan
    check = None
    if N > 1 and not a.no_check:
        ok, fnv = check_against_one_rank(sess)
        check = {"equals_one_rank_output": ok, "fnv1a64": fnv}
        if not ok:
            if ctx.is_root:
                print(json.dumps({"impl": "ours", "error": f"{N}-rank output differs from the one-rank output; "
                                  "refusing to time a wrong forward", "check": check, "gpu_launches": 0}), flush=True)
            ctx.finalize()
            sys.exit(3)
    total_ms, launches, clocks = measure_device(sess)
    if ctx.max_float(1.0 if rejected(clocks) else 0.0) > 0:   # any rank slowed down (or clock-locked): all re-measure, once
        total_ms, launches, clocks = measure_device(sess)
    clocks.pop("samples", None)
    host_binding = getattr(sess, "host_binding", None)
    del sess
    torch.cuda.empty_cache()

    alternatives = None
    if a.alt and N > 1:
        alternatives = {
            "as_timed_images_per_s": quick_ms(strategy=a.strategy, scatter=a.scatter, row_groups=rg, halo=a.halo),
            "halo_pull_images_per_s": quick_ms(strategy=a.strategy, scatter=a.scatter, row_groups=rg, halo="pull"),
            "p2p_scatter_images_per_s": quick_ms(strategy=a.strategy, scatter="p2p", row_groups=rg, halo=a.halo),
            "pure_rows_images_per_s": quick_ms(strategy=a.strategy, scatter=a.scatter, row_groups=None, halo=a.halo),
            "recompute_images_per_s": quick_ms(strategy="recompute", scatter=a.scatter, row_groups=rg),
            "note": "warm back-to-back forwards, no L2 flush (not comparable with value)"}

    e2e = None
    if not a.no_e2e:
        es = Session(cfg, w, dist=dist_arg, device=dev, precision=precision, strategy=a.strategy,
                     scatter=a.scatter, row_groups=rg, gather=a.e2e_gather, slots=4, halo=a.halo)
        e2e_ms, h2d, d2h, last = measure_e2e(es)
        h2d_all = sum(ctx.all_gather_object(h2d))
        d2h_all = sum(ctx.all_gather_object(d2h))
        e2e = {"value": round(a.batch * a.steps / (e2e_ms / 1e3), 2), "unit": UNIT,
               "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h_all,
               "strategy": strategy, "scatter": scatter, "row_groups": plan.row_groups,
               "gather": es.gather if N > 1 else "local", "steps_in_flight": es.slots - 1,
               "result_checked": bool(last is not None and torch.isfinite(last).all()) if ctx.is_root else None}
        del es
    ms_per_step = total_ms / a.steps
    value = a.batch / (ms_per_step / 1e3)
    base = REF_V4_BEST_IMG_S.get(N)
    line = {
        "metric": METRIC,
        "value": round(value, 2), "unit": UNIT, "n_gpus": N, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": HIGHER_IS_BETTER,
        "scaling": SCALING, "vs_baseline": round(value / base, 2) if base else None,
        "dtype": "tf32 (fp32 in/out, fp32 accumulate)" if precision == "tf32" else "fp32",
        "data": "synthetic (glibc rand() stream, seed 1), random-init weights",
        "config": {"model": MODEL, "implementation": "b200cnn (this repository)",
                   "timed_region": "value: CUDA events around each forward, max over ranks; e2e: wall clock over pinned "
                                   "H2D + forward + D2H of every step through Session",
                   "global_batch": a.batch, "input": f"{a.height}x{a.width}x3",
                   "parallelism": (f"spatial rows x batch: {plan.row_groups} row groups x {plan.batch_groups} batch groups"
                                   if N > 1 else "single-gpu"),
                   "strategy": strategy, "halo": a.halo if N > 1 else None, "scatter": scatter, "gather": "rank0 (P2P stores from the LRN epilogue)" if N > 1 else "local",
                   "check": check, "alternatives": alternatives, "host_binding": host_binding,
                   "vs_baseline_note": "value / the reference's published best V4 images/s at this np (other hardware, batch 1)",
                   "l2": "flushed between timed iterations (256 MiB fill, untimed)"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "impl": "ours",
    }
    if ctx.is_root:
        print(json.dumps(line), flush=True)
    ctx.finalize()


def run_reference(a):
    try:
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        import ref_arm
    except Exception as ex:  # pragma: no cover
        print(json.dumps({"impl": "reference", "unavailable": f"baseline/ref_arm.py: {ex}"}))
        return
    ref_arm.main(a)


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
