// Symmetric (peer-mapped) device memory and cross-GPU flag kernels.
//
// The reference moves every message through host std::vectors over TCP MPI
// (ref v4_mpi_cuda/src/main_mpi_cuda.cpp:58-79,124-130).  Here each rank
// cudaMalloc's a zeroed buffer, exports a CUDA IPC handle, and maps every peer's buffer into its
// own address space; kernels then ld/st peer memory directly over NVLink 5 / NVSwitch.
// torch.distributed is used only to exchange the 64-byte handles.
//
// Flag kernels implement MPI_Barrier's role (ref main_mpi_cuda.cpp:151,154) and the "gather
// complete" notification on the device, stream-ordered, without host involvement.
#include "common.cuh"

#include <cstring>

using namespace b200;

extern "C" int b200_symm_alloc(size_t bytes, void** ptr) {
    if (!ptr || bytes == 0) return B200_ERR_ARG;
    B200_CUDA_TRY(cudaMalloc(ptr, bytes));
    B200_CUDA_TRY(cudaMemset(*ptr, 0, bytes));
    B200_CUDA_TRY(cudaDeviceSynchronize());
    return B200_OK;
}

extern "C" int b200_symm_free(void* ptr) {
    if (ptr) B200_CUDA_TRY(cudaFree(ptr));
    return B200_OK;
}

extern "C" int b200_ipc_export(void* ptr, unsigned char* handle64) {
    if (!ptr || !handle64) return B200_ERR_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    B200_CUDA_TRY(cudaIpcGetMemHandle(&h, ptr));
    memcpy(handle64, &h, 64);
    return B200_OK;
}

extern "C" int b200_ipc_open(const unsigned char* handle64, void** ptr) {
    if (!ptr || !handle64) return B200_ERR_ARG;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    B200_CUDA_TRY(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return B200_OK;
}

extern "C" int b200_ipc_close(void* ptr) {
    if (ptr) B200_CUDA_TRY(cudaIpcCloseMemHandle(ptr));
    return B200_OK;
}

extern "C" int b200_memcpy2d_h2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width,
                                 size_t height, void* stream) {
    if (!dst || !src) return B200_ERR_ARG;
    if (width == 0 || height == 0) return B200_OK;
    B200_CUDA_TRY(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, height, cudaMemcpyHostToDevice,
                                    (cudaStream_t)stream));
    return B200_OK;
}

extern "C" int b200_memcpy2d_d2h(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width,
                                 size_t height, void* stream) {
    if (!dst || !src) return B200_ERR_ARG;
    if (width == 0 || height == 0) return B200_OK;
    B200_CUDA_TRY(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, height, cudaMemcpyDeviceToHost,
                                    (cudaStream_t)stream));
    return B200_OK;
}

// Host memory that several processes share (POSIX shm mapped by each of them): pin it for DMA and get the
// address kernels use for it.
extern "C" int b200_host_register(void* host, size_t bytes, void** dev_ptr) {
    if (!host || !bytes) return B200_ERR_ARG;
    cudaError_t e = cudaHostRegister(host, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
    if (e == cudaErrorHostMemoryAlreadyRegistered) { cudaGetLastError(); e = cudaSuccess; }
    B200_CUDA_TRY(e);
    if (dev_ptr) B200_CUDA_TRY(cudaHostGetDevicePointer(dev_ptr, host, 0));
    return B200_OK;
}

extern "C" int b200_host_unregister(void* host) {
    if (host && cudaHostUnregister(host) != cudaSuccess) cudaGetLastError();
    return B200_OK;
}

extern "C" int b200_can_access_peer(int dev, int peer) {
    int ok = 0;
    if (cudaDeviceCanAccessPeer(&ok, dev, peer) != cudaSuccess) { cudaGetLastError(); return 0; }
    return ok;
}

// Same-process peers (several GPUs driven by one process, or the single-GPU emulation): plain
// cudaDeviceEnablePeerAccess instead of IPC.
extern "C" int b200_enable_peer(int peer) {
    cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return B200_OK; }
    B200_CUDA_TRY(e);
    return B200_OK;
}

namespace {

struct SignalArgs { int n; int* target[16]; };

__global__ void flag_signal_kernel(SignalArgs a, int add) {
    // stream order put every earlier kernel's stores before us; publish them system-wide, then bump
    __threadfence_system();
    if ((int)threadIdx.x < a.n) red_release_sys_add(a.target[threadIdx.x], add);
}

struct WaitArgs { int n; const int* flag[16]; int target[16]; };

__global__ void flag_wait_kernel(WaitArgs a, int* status, unsigned long long timeout_ns, const int* epoch) {
    const int mult = epoch ? *epoch : 1;
    if ((int)threadIdx.x < a.n)
        wait_flag_ge(a.flag[threadIdx.x], a.target[threadIdx.x] * mult, status, timeout_ns);
    __syncthreads();
    __threadfence_system();
}

// One word of (mapped, pinned) host memory := value, after everything earlier in the stream — e.g. the D2H
// copy a reader in another process is waiting for.  A plain store: host memory takes no device atomics here.
__global__ void host_flag_store_kernel(volatile int* flag, int value) {
    __threadfence_system();
    *flag = value;
    __threadfence_system();
}

// Opens a forward: bump the local forward counter and publish "go" to the ranks whose launches are gated on
// this rank (they write into its buffers).  Everything earlier in the stream is ordered before the bump.
__global__ void flag_begin_kernel(SignalArgs a, int* epoch) {
    if (threadIdx.x == 0 && epoch) ++(*epoch);
    __threadfence_system();
    if ((int)threadIdx.x < a.n) red_release_sys_add(a.target[threadIdx.x], 1);
}

// Barrier in one launch: bump one counter on every peer, then wait for every peer's bump on ours.
// Only for ranks that own a GPU each (the waits would deadlock if ranks shared a device).
__global__ void flag_barrier_kernel(SignalArgs sig, WaitArgs w, int* status, unsigned long long timeout_ns,
                                    int* epoch) {
    __shared__ int s_epoch;
    if (threadIdx.x == 0) s_epoch = epoch ? ++(*epoch) : 0;      // this kernel opens forward number *epoch
    __syncthreads();
    __threadfence_system();
    if ((int)threadIdx.x < sig.n) red_release_sys_add(sig.target[threadIdx.x], 1);
    if ((int)threadIdx.x < w.n)
        wait_flag_ge(w.flag[threadIdx.x], epoch ? s_epoch : w.target[threadIdx.x], status, timeout_ns);
    __syncthreads();
    __threadfence_system();
}

}  // namespace

extern "C" int b200_flag_barrier(int* const* signal_targets, const int* const* wait_flags,
                                 const int* wait_target, int n, int* status, unsigned long long timeout_ns,
                                 int* epoch, void* stream) {
    if (n < 0 || n > 16 || (n && (!signal_targets || !wait_flags || !wait_target))) return B200_ERR_ARG;
    if (n == 0) return B200_OK;
    SignalArgs s;
    WaitArgs w;
    s.n = w.n = n;
    for (int i = 0; i < n; ++i) { s.target[i] = signal_targets[i]; w.flag[i] = wait_flags[i]; w.target[i] = wait_target[i]; }
    flag_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(s, w, status, timeout_ns, epoch);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_host_flag_store(int* flag_dev_ptr, int value, void* stream) {
    if (!flag_dev_ptr) return B200_ERR_ARG;
    host_flag_store_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(flag_dev_ptr, value);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_flag_begin(int* epoch, int* const* go_targets, int n, void* stream) {
    if (n < 0 || n > 16 || (n && !go_targets)) return B200_ERR_ARG;
    if (n == 0 && !epoch) return B200_OK;
    SignalArgs a;
    a.n = n;
    for (int i = 0; i < n; ++i) a.target[i] = go_targets[i];
    flag_begin_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a, epoch);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

// targets: up to 16 counters (local or peer); each gets += add
extern "C" int b200_flag_signal(int* const* targets, int n, int add, void* stream) {
    if (n < 0 || n > 16 || (n && !targets)) return B200_ERR_ARG;
    if (n == 0) return B200_OK;
    SignalArgs a;
    a.n = n;
    for (int i = 0; i < n; ++i) a.target[i] = targets[i];
    flag_signal_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a, add);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

// waits (on the stream) until flags[i] >= target[i] for all i; watchdog -> *status = TIMEOUT
extern "C" int b200_flag_wait(const int* const* flags, const int* target, int n, int* status,
                              unsigned long long timeout_ns, const int* epoch, void* stream) {
    if (n < 0 || n > 16 || (n && (!flags || !target))) return B200_ERR_ARG;
    if (n == 0) return B200_OK;
    WaitArgs a;
    a.n = n;
    for (int i = 0; i < n; ++i) { a.flag[i] = flags[i]; a.target[i] = target[i]; }
    flag_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a, status, timeout_ns, epoch);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}
