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
