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
    if (ld_acquire_sys(flag) - target >= 0) return true;
    const unsigned long long t0 = globaltimer_ns();
    while (true) {
        if (ld_acquire_sys(flag) - target >= 0) return true;
        if (status && *(volatile int*)status != 0) return false;          // someone already failed
        if (globaltimer_ns() - t0 > timeout_ns) {
            if (status) atomicCAS(status, 0, B200_ERR_TIMEOUT);
            return false;
        }
        __nanosleep(64);
    }
}
#endif

}  // namespace b200
