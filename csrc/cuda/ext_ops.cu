// Small kernels of the full-AlexNet extension (conv3-5 via im2col + tc_gemm, pool5, bias/ReLU after a
// reduce-scatter, softmax).  None of this exists in the reference, which stops after LRN2
// (planned only: ref final_project/PROBLEMS.txt:188-190, summary.md:37-43).
#include "common.cuh"

namespace {

// NHWC [N][H][W][C] -> rows [N*Ho*Wo][F*F*C], k = (fy*F+fx)*C + c, zero padding P, stride S.
// C % 4 == 0: one float4 per thread.
__global__ void im2col_nhwc(const float* __restrict__ in, float* __restrict__ col, int N, int H, int W, int C,
                            int F, int S, int P, int Ho, int Wo) {
    const int C4 = C >> 2;
    const size_t total = (size_t)N * Ho * Wo * F * F * C4;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int c4 = (int)(i % C4); size_t t = i / C4;
        const int fx = (int)(t % F); t /= F;
        const int fy = (int)(t % F); t /= F;
        const int wo = (int)(t % Wo); t /= Wo;
        const int ho = (int)(t % Ho); const int n = (int)(t / Ho);
        const int h = ho * S - P + fy, w = wo * S - P + fx;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (h >= 0 && h < H && w >= 0 && w < W)
            v = *reinterpret_cast<const float4*>(in + (((size_t)n * H + h) * W + w) * C + 4 * c4);
        reinterpret_cast<float4*>(col)[i] = v;
    }
}

__global__ void maxpool_nhwc(const float* __restrict__ in, float* __restrict__ out, int N, int H, int W, int C,
                             int F, int S, int Ho, int Wo) {
    const size_t total = (size_t)N * Ho * Wo * C;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % C); size_t t = i / C;
        const int wo = (int)(t % Wo); t /= Wo;
        const int ho = (int)(t % Ho); const int n = (int)(t / Ho);
        float m = -INFINITY;
        for (int fy = 0; fy < F; ++fy)
            for (int fx = 0; fx < F; ++fx) {
                const int h = ho * S + fy, w = wo * S + fx;
                if (h < H && w < W) m = fmaxf(m, in[(((size_t)n * H + h) * W + w) * C + c]);
            }
        out[i] = m;
    }
}

__global__ void bias_act(float* __restrict__ x, const float* __restrict__ bias, size_t rows, int cols, int ld,
                         int relu) {
    const size_t total = rows * (size_t)cols;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / cols; const int c = (int)(i % cols);
        float v = x[r * ld + c] + (bias ? bias[c] : 0.f);
        x[r * ld + c] = relu ? fmaxf(v, 0.f) : v;
    }
}

// one warp per row
__global__ void softmax_rows(const float* __restrict__ x, float* __restrict__ y, int rows, int cols, int ld) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float* xr = x + (size_t)row * ld;
    float m = -INFINITY;
    for (int c = lane; c < cols; c += 32) m = fmaxf(m, xr[c]);
    for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float s = 0.f;
    for (int c = lane; c < cols; c += 32) s += __expf(xr[c] - m);
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float inv = 1.f / s;
    for (int c = lane; c < cols; c += 32) y[(size_t)row * cols + c] = __expf(xr[c] - m) * inv;
}

// Space-to-depth by S of an NHWC row slab: out[n][Y][X][(dy*S+dx)*C + c] = in[n][S*Y+dy][S*X+dx][c],
// zeros outside the image or the slab.  Turns a stride-S FxF convolution (conv1: 11x11/4 on 3
// channels) into a stride-1 ceil(F/S)^2 convolution on S*S*C channels whose im2col rows are aligned
// 16 B chunks of the tensor itself, which the tcgen05 loader can cp.async without staging.
__global__ void space_to_depth(const float* __restrict__ in, float* __restrict__ out, int N, int H, int W, int C,
                               int S, int in_base, int in_slab, int y0, int ys, int Ws) {
    const int Cs = S * S * C;
    const size_t total = (size_t)N * ys * Ws * Cs;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int cs = (int)(i % Cs); size_t t = i / Cs;
        const int X = (int)(t % Ws); t /= Ws;
        const int Yl = (int)(t % ys); const int n = (int)(t / ys);
        const int c = cs % C, d = cs / C, dx = d % S, dy = d / S;
        const int h = S * (y0 + Yl) + dy, w = S * X + dx;
        float v = 0.f;
        if (h < H && w < W && h >= in_base && h < in_base + in_slab)
            v = in[(((size_t)n * in_slab + (h - in_base)) * W + w) * C + c];
        uint32_t u;                                   // feeds a TF32 tensor-core operand: round once here
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
        out[i] = __uint_as_float(u);
    }
}

inline unsigned nblocks(size_t n) {
    size_t b = (n + 255) / 256;
    return (unsigned)(b > 148 * 16 ? 148 * 16 : (b ? b : 1));
}

}  // namespace

extern "C" int b200_im2col_nhwc(const float* in, float* col, int N, int H, int W, int C, int F, int S, int P,
                                void* stream) {
    if (!in || !col || C % 4 != 0 || N <= 0 || H + 2 * P < F || W + 2 * P < F) return B200_ERR_ARG;
    const int Ho = (H - F + 2 * P) / S + 1, Wo = (W - F + 2 * P) / S + 1;
    im2col_nhwc<<<nblocks((size_t)N * Ho * Wo * F * F * (C / 4)), 256, 0, (cudaStream_t)stream>>>(
        in, col, N, H, W, C, F, S, P, Ho, Wo);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_maxpool_nhwc(const float* in, float* out, int N, int H, int W, int C, int F, int S,
                                 void* stream) {
    if (!in || !out || N <= 0 || H < F || W < F || S <= 0) return B200_ERR_ARG;
    const int Ho = (H - F) / S + 1, Wo = (W - F) / S + 1;
    maxpool_nhwc<<<nblocks((size_t)N * Ho * Wo * C), 256, 0, (cudaStream_t)stream>>>(in, out, N, H, W, C, F, S, Ho, Wo);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_bias_act(float* x, const float* bias, long long rows, int cols, int ld, int relu, void* stream) {
    if (!x || rows < 0 || cols <= 0 || ld < cols) return B200_ERR_ARG;
    if (rows == 0) return B200_OK;
    bias_act<<<nblocks((size_t)rows * cols), 256, 0, (cudaStream_t)stream>>>(x, bias, (size_t)rows, cols, ld, relu);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_softmax_rows(const float* x, float* y, int rows, int cols, int ld, void* stream) {
    if (!x || !y || rows <= 0 || cols <= 0 || ld < cols) return B200_ERR_ARG;
    softmax_rows<<<(rows + 7) / 8, 256, 0, (cudaStream_t)stream>>>(x, y, rows, cols, ld);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_space_to_depth(const float* in, float* out, int N, int H, int W, int C, int S, int in_base,
                                   int in_slab, int y0, int ys, void* stream) {
    if (!in || !out || N <= 0 || S <= 0 || ys < 0) return B200_ERR_ARG;
    if (ys == 0) return B200_OK;
    const int Ws = (W + S - 1) / S;
    space_to_depth<<<nblocks((size_t)N * ys * Ws * S * S * C), 256, 0, (cudaStream_t)stream>>>(
        in, out, N, H, W, C, S, in_base, in_slab, y0, ys, Ws);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}
