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

// out[r][c] = act(sum_k ws[k*stride + r*cols + c] + bias[c]), summed in k order (deterministic split-K finish)
__global__ void splitk_reduce(const float* __restrict__ ws, int ksplit, size_t stride, size_t rows, int cols,
                              const float* __restrict__ bias, float* __restrict__ out, int ld, int flags) {
    const size_t total = rows * (size_t)cols;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / cols; const int c = (int)(i % cols);
        float v = 0.f;
        for (int k = 0; k < ksplit; ++k) v += ws[k * stride + i];
        if (flags & 8) v += bias[c];
        if (flags & 1) v = fmaxf(v, 0.f);
        if (flags & 2) {
            unsigned u;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
            v = __uint_as_float(u);
        }
        out[r * ld + c] = v;
    }
}

// out[r][c] = act(sum_k parts[k][r*ld_in + c] + bias[c]), summed in k order.  parts[k] may be peer-mapped pointers:
// the owner of a shard PULLS every rank's partial sums over NVLink and adds them in rank order — a reduce-scatter
// that is deterministic (no atomics, fixed order) and needs no zeroed accumulator.
struct PeerParts { const float* p[16]; int n; };
__global__ void peer_reduce(PeerParts parts, size_t rows, int cols, int ld_in, const float* __restrict__ bias,
                            float* __restrict__ out, int ld, int flags) {
    const int c4n = cols >> 2;
    const size_t total = rows * (size_t)c4n;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / c4n; const int c = (int)(i % c4n) * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int k = 0; k < parts.n; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(parts.p[k] + r * ld_in + c);
            v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
        }
        float o[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (flags & 8) o[j] += bias[c + j];
            if (flags & 1) o[j] = fmaxf(o[j], 0.f);
            if (flags & 2) {
                unsigned u;
                asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(o[j]));
                o[j] = __uint_as_float(u);
            }
        }
        *reinterpret_cast<float4*>(out + r * ld + c) = make_float4(o[0], o[1], o[2], o[3]);
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
    // One block per (image, s2d row): the S input rows are read coalesced into shared memory, then the
    // s2d row (Ws pixels x S*S*C floats, contiguous) is written coalesced — whole 128 B lines both ways.
    extern __shared__ float tile[];                               // [S][rowf]
    const int SC = S * C, Cs = S * SC, rowf = Ws * SC, WC = W * C;
    const int nrows = N * ys;
    for (int row = blockIdx.x; row < nrows; row += gridDim.x) {
        const int Yl = row % ys, n = row / ys;
        for (int i = threadIdx.x; i < S * rowf; i += blockDim.x) {
            const int dy = i / rowf, f = i - dy * rowf;
            const int h = S * (y0 + Yl) + dy;
            float v = 0.f;
            if (h < H && h >= in_base && h < in_base + in_slab && f < WC)
                v = in[((size_t)n * in_slab + (h - in_base)) * WC + f];
            uint32_t u;                               // feeds a TF32 tensor-core operand: round once here
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
            tile[i] = __uint_as_float(u);
        }
        __syncthreads();
        float* dst = out + (size_t)row * Ws * Cs;
        for (int o = threadIdx.x; o < Ws * Cs; o += blockDim.x) {
            const int X = o / Cs, cs = o - X * Cs;
            const int dy = cs / SC, r = cs - dy * SC;
            dst[o] = tile[dy * rowf + X * SC + r];
        }
        __syncthreads();
    }
}

// Fast path with compile-time S and C (conv1: S=4, C=3): constant divisions, 128-bit smem reads and
// global writes, one block per (image, s2d row).
template <int S, int C>
__global__ void __launch_bounds__(256)
space_to_depth_fixed(const float* __restrict__ in, float* __restrict__ out, int H, int W, int in_base, int in_slab,
                     int y0, int ys, int Ws) {
    constexpr int SC = S * C, Cs = S * SC;
    static_assert(SC % 4 == 0, "a dy-run must be whole float4s");
    extern __shared__ __align__(16) float tile[];                 // [S][rowf]
    const int rowf = Ws * SC, WC = W * C;
    const int row = blockIdx.x, Yl = row % ys, n = row / ys;
#pragma unroll
    for (int dy = 0; dy < S; ++dy) {
        const int h = S * (y0 + Yl) + dy;
        const bool rv = h < H && h >= in_base && h < in_base + in_slab;
        const float* src = in + ((size_t)n * in_slab + (rv ? h - in_base : 0)) * WC;
        for (int f = threadIdx.x; f < rowf; f += 256) {
            const float v = (rv && f < WC) ? __ldg(src + f) : 0.f;
            uint32_t u;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
            tile[dy * rowf + f] = __uint_as_float(u);
        }
    }
    __syncthreads();
    float4* dst = reinterpret_cast<float4*>(out + (size_t)row * Ws * Cs);
    constexpr int Q = Cs / 4, QR = SC / 4;                        // float4 per pixel / per dy-run
    for (int o = threadIdx.x; o < Ws * Q; o += 256) {
        const int X = o / Q, q = o - X * Q;
        const int dy = q / QR, r4 = q - dy * QR;
        dst[o] = *reinterpret_cast<const float4*>(tile + dy * rowf + X * SC + r4 * 4);
    }
}

// contiguous rows (width floats) -> rows at a larger pitch; coalesced both ways
__global__ void repitch_rows(const float* __restrict__ src, float* __restrict__ dst, long long rows, int width,
                             int dpitch) {
    for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
        const float* s = src + r * width;
        float* d = dst + r * dpitch;
        for (int f = blockIdx.x * blockDim.x + threadIdx.x; f < width; f += gridDim.x * blockDim.x) d[f] = s[f];
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

// finish of b200_tc_gemm_splitk: ws holds ksplit partial [rows][cols] matrices `stride` floats apart.
// flags: 1 ReLU, 2 round to TF32, 8 add bias (same bits as b200_tc_gemm).
extern "C" int b200_splitk_reduce(const float* ws, int ksplit, long long stride, long long rows, int cols,
                                  const float* bias, float* out, int ld, int flags, void* stream) {
    if (!ws || !out || ksplit < 1 || rows < 0 || cols <= 0 || ld < cols || stride < rows * cols) return B200_ERR_ARG;
    if ((flags & 8) && !bias) return B200_ERR_ARG;
    if (rows == 0) return B200_OK;
    splitk_reduce<<<nblocks((size_t)rows * cols), 256, 0, (cudaStream_t)stream>>>(ws, ksplit, (size_t)stride,
                                                                               (size_t)rows, cols, bias, out, ld, flags);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

// flags: 1 ReLU, 2 round to TF32, 8 add bias (same bits as b200_tc_gemm); cols, ld_in and ld multiples of 4
extern "C" int b200_peer_reduce(const float* const* parts, int n, long long rows, int cols, int ld_in,
                                const float* bias, float* out, int ld, int flags, void* stream) {
    if (!parts || !out || n < 1 || n > 16 || rows < 0 || cols <= 0 || (cols & 3) || (ld_in & 3) || (ld & 3) ||
        ld_in < cols || ld < cols)
        return B200_ERR_ARG;
    if ((flags & 8) && !bias) return B200_ERR_ARG;
    if (rows == 0) return B200_OK;
    PeerParts pp;
    pp.n = n;
    for (int i = 0; i < n; ++i) { if (!parts[i]) return B200_ERR_ARG; pp.p[i] = parts[i]; }
    peer_reduce<<<nblocks((size_t)rows * (cols / 4)), 256, 0, (cudaStream_t)stream>>>(pp, (size_t)rows, cols, ld_in, bias,
                                                                                    out, ld, flags);
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
    const int rowf = Ws * S * C, nrows = N * ys;
    const size_t smem = (size_t)S * rowf * sizeof(float);
    if (S == 4 && C == 3 && smem <= 48 * 1024) {
        space_to_depth_fixed<4, 3><<<nrows, 256, smem, (cudaStream_t)stream>>>(in, out, H, W, in_base, in_slab,
                                                                             y0, ys, Ws);
        B200_CUDA_TRY(cudaGetLastError());
        return B200_OK;
    }
    if (smem > 200 * 1024) return B200_ERR_UNSUPPORTED;
    B200_CUDA_TRY(cudaFuncSetAttribute(space_to_depth, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    space_to_depth<<<nrows < 148 * 8 ? nrows : 148 * 8, 256, smem, (cudaStream_t)stream>>>(
        in, out, N, H, W, C, S, in_base, in_slab, y0, ys, Ws);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_repitch_rows(const float* src, float* dst, long long rows, int width, int dpitch, void* stream) {
    if (!src || !dst || rows < 0 || width <= 0 || dpitch < width) return B200_ERR_ARG;
    if (rows == 0) return B200_OK;
    dim3 grid((width + 255) / 256, (unsigned)(rows < 65535 ? rows : 65535));
    repitch_rows<<<grid, 256, 0, (cudaStream_t)stream>>>(src, dst, rows, width, dpitch);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}
