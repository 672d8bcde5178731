"""Exhaustive check of the barrier-free multi-GPU forward protocol (models/engine.py::_enqueue_stages).

compute-sanitizer's racecheck cannot see races BETWEEN GPUs, and it is closed on this pool anyway, so the argument
that the forward needs no barrier is checked here on a model instead: every rank is a stream of stages

    begin(k)   ++epoch; rank 0 also publishes go = k to every rank
    block1(k)  writes its own pool1 rows into slab parity k % 2 and (halo push) the boundary rows into each
               neighbour's slab of the same parity, then bumps the neighbour's arrival counter
    block2(k)  gated on go >= k; waits for the neighbours' arrival counters; reads the slab (own rows + halo rows;
               halo pull: the neighbours' slabs instead); stores its output rows; its end bumps rank 0's done counter
    wait(k)    rank 0 only: until every rank's done counter has reached k

plus the consumer of the results on rank 0 (the host, or a copy stream): it reads output slot k % 2 after wait(k), and
— the caller's side of the contract — rank 0 does not begin forward k + 2 before result k has been consumed.  All of it
for k = 1..K, executed in stream order per rank and in ANY interleaving across ranks (a stage blocks only on the
flags it really waits for).  block1 and block2 are intervals (start / end events), so "a kernel is still reading
while another one writes" is representable.  The search visits every reachable state and reports

* a hazard: a region written by forward k while forward j != k is still being read or written there, or a region
  read for forward k whose last completed write is not forward k (stale or torn data);
* a deadlock: some rank is not finished and nobody can move.

The reference has neither the problem nor the check: its ranks meet in blocking MPI calls
(ref final_project/v4_mpi_cuda/src/main_mpi_cuda.cpp:61-62,127-130) and in MPI_Barrier.  ``mutations`` remove one
ingredient of the protocol (the go gate, the parity double buffer, the arrival wait, rank 0's completion wait) so
the tests can show that the checker does find the race each ingredient is there to prevent.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, FrozenSet, List, Optional, Tuple

BEGIN, B1S, B1E, B2S, B2E, WAIT, CS, CE = range(8)
_NAMES = {BEGIN: "begin", B1S: "block1.start", B1E: "block1.end", B2S: "block2.start", B2E: "block2.end", WAIT: "wait",
          CS: "consume.start", CE: "consume.end"}


@dataclass(frozen=True)
class Finding:
    kind: str                 # "hazard" | "deadlock"
    detail: str
    trace: Tuple[str, ...]    # the interleaving that leads to it


def neighbours(rank: int, world: int) -> List[int]:
    """Row decomposition: a rank exchanges halo rows with the ranks next to it."""
    return [r for r in (rank - 1, rank + 1) if 0 <= r < world]


def program(rank: int, forwards: int) -> List[Tuple[int, int]]:
    ops = []
    for k in range(1, forwards + 1):
        ops += [(BEGIN, k), (B1S, k), (B1E, k), (B2S, k), (B2E, k)]
        if rank == 0:
            ops.append((WAIT, k))
    return ops


def check(world: int = 2, forwards: int = 4, halo: str = "push", mutations: FrozenSet[str] = frozenset(),
          max_states: int = 2_000_000) -> Optional[Finding]:
    """Explore every interleaving; returns the first Finding or None when the protocol is safe for this shape.

    mutations: "no_go_gate" (block 2 does not wait for go >= k), "single_slab" (one pool1 slab instead of two),
    "no_arrival_wait" (block 2 does not wait for the neighbours' counters), "no_done_wait" (rank 0 publishes the
    next go without waiting for the ranks' completion counters)."""
    assert halo in ("push", "pull")
    progs = [program(r, forwards) for r in range(world)]
    progs.append([ev for k in range(1, forwards + 1) for ev in ((CS, k), (CE, k))])     # actor `world`: the consumer
    CONS = world
    nbrs = [neighbours(r, world) for r in range(world)]
    par = (lambda k: 0) if "single_slab" in mutations else (lambda k: k & 1)

    def counters(pcs):
        """Flag values implied by how far every rank has got (flags only grow, so they are functions of the pcs)."""
        go = waited = consumed = 0
        done = [0] * world
        b1_done = [0] * world              # last forward whose block 1 (stores, pushes, counter bumps) has finished
        for r, pc in enumerate(pcs):
            for op, k in progs[r][:pc]:
                if op == BEGIN and r == 0:
                    go = k
                elif op == B1E:
                    b1_done[r] = k
                elif op == B2E:
                    done[r] = k
                elif op == WAIT:
                    waited = k
                elif op == CE:
                    consumed = k
        return go, done, b1_done, waited, consumed

    def active(pcs):
        """(rank, 'block1'|'block2', k) of every kernel that has started and not ended."""
        out = []
        for r, pc in enumerate(pcs):
            if pc == 0:
                continue
            op, k = progs[r][pc - 1]
            if op == B1S:
                out.append((r, "block1", k))
            elif op == B2S:
                out.append((r, "block2", k))
            elif op == CS:
                out.append((r, "consume", k))
        return out

    def regions(r, stage, k):
        """(writes, reads) of a running kernel: region = (owner rank of the slab, parity, part)."""
        p = par(k)
        if stage == "consume":                 # the gathered result: every rank's rows of output slot k % 2 on rank 0
            return [], [(0, k & 1, f"out_rows_of_{q}") for q in range(world)]
        if stage == "block1":
            w = [(r, p, "own")]
            if halo == "push":
                w += [(n, p, f"halo_from_{r}") for n in nbrs[r]]
            return w, []
        rd = [(r, p, "own")]
        if halo == "push":
            rd += [(r, p, f"halo_from_{n}") for n in nbrs[r]]
        else:
            rd += [(n, p, "own") for n in nbrs[r]]
        return [(0, k & 1, f"out_rows_of_{r}")], rd

    def hazard(pcs, last_write) -> Optional[str]:
        act = [("the consumer's" if r == CONS else f"rank {r}", st, k, regions(r, st, k)) for r, st, k in active(pcs)]
        for i, (who, st, k, (w, rd)) in enumerate(act):
            for (who2, st2, k2, (w2, rd2)) in act[i + 1:]:      # two running kernels, one writing where the other works
                for reg in w:
                    if reg in w2 or reg in rd2:
                        return (f"{who} {st}({k}) writes region {reg} while {who2} {st2}({k2}) "
                                f"{'writes' if reg in w2 else 'reads'} it")
                for reg in rd:
                    if reg in w2:
                        return f"{who2} {st2}({k2}) writes region {reg} while {who} {st}({k}) reads it"
            for reg in rd:                                      # the data a reader sees must be this forward's
                if last_write.get(reg) != k:
                    return (f"{who} {st}({k}) reads region {reg} whose last completed write is forward "
                            f"{last_write.get(reg)}")
        return None

    def enabled(r, pcs, go, done, b1_done, waited, consumed) -> bool:
        op, k = progs[r][pcs[r]]
        if op == CS:
            return waited >= k                 # the result is handed to the consumer after rank 0's completion wait
        if op == BEGIN and r == 0 and k > 2 and consumed < k - 2:
            return False                       # caller's contract: forward k reuses the slot of result k - 2
        if op == B2S:
            if "no_go_gate" not in mutations and go < k:
                return False
            if "no_arrival_wait" not in mutations and any(b1_done[n] < k for n in nbrs[r]):
                return False
        if op == WAIT and "no_done_wait" not in mutations and any(d < k for d in done):
            return False
        return True

    start = tuple([0] * (world + 1))
    # last completed write per region depends on the path only through the pcs (writes complete in program order per
    # region writer), so it is recomputed from the pcs too
    def last_writes(pcs) -> Dict[tuple, int]:
        lw: Dict[tuple, int] = {}
        for r, pc in enumerate(pcs):
            for op, k in progs[r][:pc]:
                if op in (B1E, B2E):
                    for reg in regions(r, "block1" if op == B1E else "block2", k)[0]:
                        # two ranks never write the same region, and one rank's kernels are in stream order
                        lw[reg] = k
        return lw

    seen = {start}
    stack: List[Tuple[tuple, Tuple[str, ...]]] = [(start, ())]
    while stack:
        pcs, trace = stack.pop()
        go, done, b1_done, waited, consumed = counters(pcs)
        moved = False
        unfinished = [r for r in range(world + 1) if pcs[r] < len(progs[r])]
        for r in unfinished:
            if not enabled(r, pcs, go, done, b1_done, waited, consumed):
                continue
            moved = True
            nxt = list(pcs)
            nxt[r] += 1
            nxt = tuple(nxt)
            if nxt in seen:
                continue
            seen.add(nxt)
            if len(seen) > max_states:
                raise RuntimeError("state space larger than max_states")
            op, k = progs[r][pcs[r]]
            t2 = trace + ((f"r{r}" if r != CONS else "consumer") + f":{_NAMES[op]}({k})",)
            h = hazard(nxt, last_writes(nxt))
            if h:
                return Finding("hazard", h, t2)
            stack.append((nxt, t2))
        if unfinished and not moved:
            return Finding("deadlock", f"ranks {unfinished} cannot move at {pcs}", trace)
    return None
