// Shared host/device helpers for the b200cnn CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include "b200cnn.h"

namespace b200 {

// ---- host error plumbing (the reference's CUDA_CHECK prints and exits / MPI_Aborts,
// ref v3_cuda_only/src/alexnet_cuda.cu:10-19, v4_mpi_cuda/src/layers_mpi_cuda.cu:14-20; a library
// must not kill its caller, so the message is latched and a status is returned) ---------------
void set_error(const char* fmt, ...);

#define B200_CUDA_TRY(expr)                                                              \
    do {                                                                                 \
        cudaError_t e__ = (expr);                                                        \
        if (e__ != cudaSuccess) {                                                        \
            ::b200::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr,               \
                              cudaGetErrorString(e__));                                  \
            return B200_ERR_CUDA;                                                        \
        }                                                                                \
    } while (0)

// ---- device: cross-GPU flag protocol ------------------------------------------------------
// Counters live in peer-mapped ("symmetric") memory and only ever grow; a consumer waits for
// `flag >= target`.  Producer: data stores -> bar.sync -> one thread: fence.sys + red.release.sys.
// Consumer: one thread spins on ld.acquire.sys -> bar.sync -> everybody loads.
// The watchdog turns a dead peer into an error code instead of a hung GPU.
#ifdef __CUDACC__
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_sys_add(int* p, int v) {
    asm volatile("red.release.sys.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// returns true when the flag reached the target, false on watchdog expiry (status latched)
__device__ __forceinline__ bool wait_flag_ge(const int* flag, int target, int* status,
                                             unsigned long long timeout_ns) {
    unsigned spins = 0;
    unsigned long long t0 = 0;
    while (true) {
        if (ld_acquire_sys(flag) - target >= 0) return true;
        // the first polls are back to back (a flag that is about to flip is seen at load latency); the error latch
        // and the timer — the expensive reads — are looked at every 64th poll only
        if ((++spins & 63) == 0) {
            if (status && *(volatile int*)status != 0) return false;      // someone already failed
            const unsigned long long now = globaltimer_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > timeout_ns) {
                if (status) atomicCAS(status, 0, B200_ERR_TIMEOUT);
                return false;
            }
        }
        if (spins > 32) __nanosleep(32);
    }
}
// Arrival signals of one finished tile that produced pooled rows [r0, r1): called by ONE thread after a barrier
// that follows the tile's last store (local and pushed).  red.release.sys orders those stores (made visible to
// this thread by the barrier) before the bump at system scope; no separate fence is needed.
__device__ __forceinline__ bool tile_signals(const b200_block_geom& g, int r0, int r1) {
    for (int i = 0; i < g.n_signal; ++i)
        if (g.signal_hi[i] <= g.signal_lo[i] || (r0 < g.signal_hi[i] && r1 > g.signal_lo[i])) return true;
    return false;
}
__device__ __forceinline__ void tile_signal(const b200_block_geom& g, int n, int r0, int r1) {
    for (int i = 0; i < g.n_signal; ++i)
        if (g.signal_hi[i] <= g.signal_lo[i] || (r0 < g.signal_hi[i] && r1 > g.signal_lo[i]))
            red_release_sys_add(g.signal[i] + n, 1);
}

// Flow-control gate and completion notice of a block launch (b200_block_geom::gate / done*).
// gate: called by ONE thread per CTA before the CTA-wide barrier that ends the kernel prologue.
__device__ __forceinline__ void block_gate_wait(const b200_block_geom& g) {
    if (g.gate) wait_flag_ge(g.gate, g.gate_target * (g.epoch ? *g.epoch : 1), g.status, g.timeout_ns);
}
// done: called by ONE thread per CTA after a CTA-wide barrier that follows the CTA's last store.  The
// fences order every CTA's stores (made visible to this thread by that barrier) before the notice.
__device__ __forceinline__ void block_done_signal(const b200_block_geom& g) {
    if (!g.done_count) return;
    __threadfence_system();
    const unsigned int old = atomicAdd(g.done_count, 1u);
    if ((old + 1u) % gridDim.x == 0u) {
        __threadfence_system();
        for (int i = 0; i < g.n_done; ++i) red_release_sys_add(g.done[i], 1);
    }
}
#endif

}  // namespace b200
