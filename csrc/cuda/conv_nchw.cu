// Stand-alone convolution on NCHW tensors with reference-layout weights [K][C][F][F]: the
// framework-side equivalent of the reference's `conv_test` program (final_project/conv_test, a
// binary whose sources are not in the tree; its kernel convTiledKernel is NCHW).
//
// The framework's native layout is NHWC with [F][F][C][K] weights, so this entry point is an
// adaptor: it transposes the input, re-packs the weights, runs the NHWC direct convolution and
// transposes the result back.  Those three intermediates are SCRATCH — device memory beyond the
// caller's inputs and outputs:
//     [ NHWC input | packed weights | NHWC output ]      (b200_conv_nchw_scratch_bytes)
// b200_conv_nchw allocates the scratch on the stream for the duration of the call.
#include "common.cuh"

#include <mutex>

namespace {

__global__ void nchw_to_nhwc(const float* __restrict__ in, float* __restrict__ out, int N, int C, int H, int W) {
    const size_t total = (size_t)N * C * H * W;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % C); size_t t = i / C;
        const int w = (int)(t % W); t /= W;
        const int h = (int)(t % H); const int n = (int)(t / H);
        out[i] = in[(((size_t)n * C + c) * H + h) * W + w];
    }
}

__global__ void nhwc_to_nchw(const float* __restrict__ in, float* __restrict__ out, int N, int C, int H, int W) {
    const size_t total = (size_t)N * C * H * W;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int w = (int)(i % W); size_t t = i / W;
        const int h = (int)(t % H); t /= H;
        const int c = (int)(t % C); const int n = (int)(t / C);
        out[i] = in[(((size_t)n * H + h) * W + w) * C + c];
    }
}

__global__ void pack_kcff_to_ffck(const float* __restrict__ w, float* __restrict__ o, int K, int C, int F) {
    const int total = K * C * F * F;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int k = i % K; int t = i / K;
        const int c = t % C; t /= C;
        const int fx = t % F; const int fy = t / F;
        o[i] = w[((k * C + c) * F + fy) * F + fx];
    }
}

// direct convolution, NHWC in / [F][F][C][K] weights / NHWC out; thread = (pixel, k), k fastest:
// weight reads are coalesced, the input value is a warp broadcast.
__global__ void conv_nhwc_plain(const float* __restrict__ in, const float* __restrict__ wt,
                                const float* __restrict__ bias, float* __restrict__ out, int N, int H, int W,
                                int C, int K, int F, int S, int P, int Ho, int Wo) {
    const size_t total = (size_t)N * Ho * Wo * K;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int k = (int)(i % K); size_t t = i / K;
        const int wo = (int)(t % Wo); t /= Wo;
        const int ho = (int)(t % Ho); const int n = (int)(t / Ho);
        float acc = bias ? bias[k] : 0.f;
        for (int fy = 0; fy < F; ++fy) {
            const int hi = ho * S - P + fy;
            if (hi < 0 || hi >= H) continue;
            for (int fx = 0; fx < F; ++fx) {
                const int wi = wo * S - P + fx;
                if (wi < 0 || wi >= W) continue;
                const float* ip = in + (((size_t)n * H + hi) * W + wi) * C;
                const float* wp = wt + ((size_t)(fy * F + fx) * C) * K + k;
                for (int c = 0; c < C; ++c) acc = fmaf(ip[c], wp[(size_t)c * K], acc);
            }
        }
        out[i] = acc;
    }
}

struct ConvDims { int N, C, H, W, K, F, S, P, Ho, Wo; size_t in_elems, w_elems, out_elems; };

bool make_dims(ConvDims* d, int N, int C, int H, int W, int K, int F, int S, int P) {
    if (N <= 0 || C <= 0 || H <= 0 || W <= 0 || K <= 0 || F <= 0 || S <= 0 || P < 0) return false;
    if (H + 2 * P < F || W + 2 * P < F) return false;
    d->N = N; d->C = C; d->H = H; d->W = W; d->K = K; d->F = F; d->S = S; d->P = P;
    d->Ho = (H - F + 2 * P) / S + 1; d->Wo = (W - F + 2 * P) / S + 1;
    d->in_elems = (size_t)N * C * H * W;
    d->w_elems = (size_t)K * C * F * F;
    d->out_elems = (size_t)N * K * d->Ho * d->Wo;
    return true;
}

inline unsigned blocks_for(size_t n) {
    size_t b = (n + 255) / 256;
    return (unsigned)(b > 148 * 32 ? 148 * 32 : (b ? b : 1));
}

// the whole adaptor pipeline on `stream`, with caller-provided scratch of >= scratch_elems floats
int run_with_scratch(const ConvDims& d, const float* in_nchw, const float* w_kcff, const float* bias,
                     float* out_nchw, float* scratch, cudaStream_t st) {
    float* s_in = scratch;
    float* s_w = s_in + d.in_elems;
    float* s_out = s_w + d.w_elems;
    nchw_to_nhwc<<<blocks_for(d.in_elems), 256, 0, st>>>(in_nchw, s_in, d.N, d.C, d.H, d.W);
    pack_kcff_to_ffck<<<blocks_for(d.w_elems), 256, 0, st>>>(w_kcff, s_w, d.K, d.C, d.F);
    conv_nhwc_plain<<<blocks_for(d.out_elems), 256, 0, st>>>(s_in, s_w, bias, s_out, d.N, d.H, d.W, d.C, d.K,
                                                             d.F, d.S, d.P, d.Ho, d.Wo);
    nhwc_to_nchw<<<blocks_for(d.out_elems), 256, 0, st>>>(s_out, out_nchw, d.N, d.K, d.Ho, d.Wo);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

}  // namespace

extern "C" long long b200_conv_nchw_scratch_bytes(int N, int C, int H, int W, int K, int F, int S, int P) {
    ConvDims d;
    if (!make_dims(&d, N, C, H, W, K, F, S, P)) return B200_ERR_ARG;
    return (long long)((d.in_elems + d.w_elems + d.out_elems) * sizeof(float));
}

extern "C" int b200_conv_nchw(const float* in_nchw, const float* w_kcff, const float* bias, float* out_nchw,
                              int N, int C, int H, int W, int K, int F, int S, int P, void* stream) {
    ConvDims d;
    if (!in_nchw || !w_kcff || !out_nchw || !make_dims(&d, N, C, H, W, K, F, S, P)) return B200_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    float* scratch = nullptr;
    B200_CUDA_TRY(cudaMallocAsync((void**)&scratch, (d.in_elems + d.w_elems + d.out_elems) * sizeof(float), st));
    int rc = run_with_scratch(d, in_nchw, w_kcff, bias, out_nchw, scratch, st);
    B200_CUDA_TRY(cudaFreeAsync(scratch, st));
    return rc;
}

// ---------------------------------------------------------------------------------------------
// Allocation-free variant.
//
// Who needs it: hosts that may not allocate device memory when they call us — an inference server
// whose own pool has already reserved the whole GPU (a cudaMalloc here would fail or fragment it),
// or a latency-bound loop that cannot afford the device-wide synchronisation cudaMalloc/cudaFree
// imply.  b200_conv_nchw (above) needs scratch for its NHWC/packed intermediates and gets it from
// cudaMallocAsync; this entry point runs the very same kernels but takes the scratch from a fixed
// array that is part of the library image.
//
// Cost and limits.  The array is B200_CONV_NCHW_STATIC_SCRATCH_BYTES (8 MiB) and every GPU that
// loads this library carries it until the process exits, used or not.  A problem whose scratch
// (b200_conv_nchw_scratch_bytes) exceeds it is refused with B200_ERR_RANGE and a message — there is
// no fallback to allocation.  8 MiB covers both AlexNet convolutions of this project for one image
// (conv1: 0.62 + 0.14 + 1.16 MB, conv2: 0.28 + 2.46 + 0.75 MB).
//
// Sharing.  There is one array per GPU and every call on that GPU uses it.  It still holds the
// previous call's data when the next call starts; that is harmless here because each call writes
// every scratch element before reading it (nothing relies on zeros), but two calls that overlap on
// the same GPU — different streams, or different host threads — would overwrite each other's
// intermediates: callers must order them (same stream, or an event between them).
//
// Set-up.  Host code cannot use the array's name as a pointer (on the host that name is only a
// handle for the runtime).  The device address is looked up with cudaGetSymbolAddress the first
// time the entry point is called on a given device, with that device current, and cached per
// device.  That first call therefore does a little extra runtime work; make it before any stream
// capture you intend to do around later calls.
//
// When the 8 MiB come into being.  The CUDA runtime loads a module's code and its __device__ data
// on a GPU lazily: on the first kernel launch or symbol lookup from that module on that GPU.  So the
// array is NOT allocated by this entry point, but the first call into this file on a device may be
// the moment the driver sets the module (array included) up there.  A host whose pool is about to
// take all remaining memory must therefore make ONE warm-up call per device (any call into this
// file: this entry point or b200_conv_nchw) BEFORE its pool reserves the GPU; after that, calls here
// neither allocate nor free anything (tests/test_zz_no_alloc.py checks the free-memory figure after
// such a warm-up).  The cached address does not survive cudaDeviceReset(): a process that resets a
// device must not call this entry point on it again.
#define B200_CONV_NCHW_STATIC_SCRATCH_BYTES (8u << 20)

static __device__ __align__(256) unsigned char g_conv_nchw_static_scratch[B200_CONV_NCHW_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
float* g_static_scratch_addr[kMaxDevices] = {};      // per device, resolved on first use
std::mutex g_static_scratch_mu;

int static_scratch_for_current_device(float** out) {
    int dev = -1;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        b200::set_error("b200_conv_nchw_noalloc: device ordinal %d outside [0,%d)", dev, kMaxDevices);
        return B200_ERR_ARG;
    }
    std::lock_guard<std::mutex> lk(g_static_scratch_mu);
    if (!g_static_scratch_addr[dev]) {
        void* p = nullptr;
        B200_CUDA_TRY(cudaGetSymbolAddress(&p, g_conv_nchw_static_scratch));   // device `dev` is current
        g_static_scratch_addr[dev] = static_cast<float*>(p);
    }
    *out = g_static_scratch_addr[dev];
    return B200_OK;
}
}  // namespace

extern "C" long long b200_conv_nchw_noalloc_max_scratch_bytes(void) {
    return (long long)B200_CONV_NCHW_STATIC_SCRATCH_BYTES;
}

extern "C" int b200_conv_nchw_noalloc(const float* in_nchw, const float* w_kcff, const float* bias,
                                      float* out_nchw, int N, int C, int H, int W, int K, int F, int S,
                                      int P, void* stream) {
    ConvDims d;
    if (!in_nchw || !w_kcff || !out_nchw || !make_dims(&d, N, C, H, W, K, F, S, P)) return B200_ERR_ARG;
    const size_t need = (d.in_elems + d.w_elems + d.out_elems) * sizeof(float);
    if (need > B200_CONV_NCHW_STATIC_SCRATCH_BYTES) {
        b200::set_error("b200_conv_nchw_noalloc: this problem needs %zu bytes of scratch but the built-in "
                        "scratch array holds %u bytes; use b200_conv_nchw (which allocates) or a smaller batch",
                        need, (unsigned)B200_CONV_NCHW_STATIC_SCRATCH_BYTES);
        return B200_ERR_RANGE;
    }
    float* scratch = nullptr;
    int rc = static_scratch_for_current_device(&scratch);
    if (rc != B200_OK) return rc;
    return run_with_scratch(d, in_nchw, w_kcff, bias, out_nchw, scratch, (cudaStream_t)stream);
}
