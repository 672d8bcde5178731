// tcgen05 GEMM for the full-AlexNet extension: C[M][N] (+)= act(A[M][K] * W[N][K]^T + bias).
//
// The reference stops after LRN2; FC6-8 (and conv3-5) are only planned there
// (ref final_project/PROBLEMS.txt:188-190, summary.md:37-43,68).  This kernel is the building block
// of that extension and of its tensor-parallel form:
//   * A is gathered by 4 producer warps with cp.async straight into the 128B-swizzled stage.  A is a
//     plain pointer, so it may be another GPU's activation buffer mapped over NVLink: the
//     all-gather of "all-gather -> GEMM" is the loader itself.
//   * W (K-major, TF32-rounded once) arrives by TMA; tcgen05.mma kind::tf32 M128 x N128 x K8,
//     fp32 accumulators double-buffered in TMEM so tile t's epilogue overlaps tile t+1's main loop.
//   * Epilogue: tcgen05.ld -> +bias -> ReLU -> either plain stores, or red.global.add.f32 into C, which
//     may be a peer pointer: the reduce(-scatter) of "GEMM -> reduce-scatter" is the epilogue itself.
// Warps: 0-3 A producers, 4 MMA issuer + TMEM owner, 5 TMA issuer, 6-9 epilogue (TMEM quarter = warp%4).
#include "tc_ptx.cuh"

#include <cstring>
#include <vector>

namespace {

using namespace b200;
using namespace b200::tc;

constexpr int kBM = 128, kBN = 128, kBK = 32;
constexpr int kStages = 5;
constexpr int kAStage = kBM * 128, kBStage = kBN * 128;          // bytes
constexpr int kGemmThreads = 320;
constexpr int kProdThreads = 128;

struct GemmParams {
    const float* A;
    float* C;
    const float* bias;
    int M, N, K, lda, ldc;
    int flags;                                                   // 1 relu, 2 round to tf32, 4 atomic add, 8 bias
    int tiles_m, tiles_n, nkb;
    uint32_t idesc;
    // implicit-GEMM convolution (kConv): A is an NHWC activation [cN][cH][cW][cC], row m of the GEMM is
    // output pixel m = (n*cHo + oy)*cWo + ox and k-block kb is 32 channels of filter tap (fy, fx):
    // kb = (fy*cF + fx)*(cC/32) + chunk.  Taps outside the image are zero-filled by cp.async (size 0).
    int cH, cW, cC, cF, cS, cP, cHo, cWo;
};

// kConv = false: A[M][lda] row-major.  kConv = true: A rows are gathered from the NHWC activation, so the
// im2col matrix of conv3-5 never exists in memory.
template <bool kConv>
__global__ void __launch_bounds__(kGemmThreads, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tmapW, const __grid_constant__ GemmParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t b_off = kStages * kAStage;
    const uint32_t ctl = b_off + kStages * kBStage;
    const uint32_t full_bar = base + ctl, empty_bar = full_bar + 8 * kStages;
    const uint32_t tfull_bar = empty_bar + 8 * kStages, tempty_bar = tfull_bar + 16;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gbase + ctl + 16 * kStages + 32);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(full_bar + 8 * s, kProdThreads + 1);
            mbar_init(empty_bar + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(tfull_bar + 8 * b, 1);
            mbar_init(tempty_bar + 8 * b, 4);
        }
        fence_mbar_init();
    }
    if (warp == 4) tmem_alloc(smem_u32(tmem_slot), 2 * kBN);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const long long total = (long long)p.tiles_m * p.tiles_n;

    if (warp < 4) {
        // ------------------------------------------------ A producers (local or peer memory)
        const int tid = threadIdx.x, j = tid & 7, r0 = tid >> 3;          // rows r0 + 16*i
        int stage = 0;
        uint32_t phase = 0;
        for (long long t = blockIdx.x; t < total; t += gridDim.x) {
            const int m0 = (int)(t / p.tiles_n) * kBM;
            // kConv: per-row pixel origin (float offset of input (n, oy*S-P, ox*S-P, 4j)) and tap validity
            long long poff[8];
            uint32_t ymask[8], xmask[8];
            if (kConv) {
                const int hw = p.cHo * p.cWo;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int m = m0 + r0 + 16 * i;
                    const int mm = m < p.M ? m : 0;
                    const int n = mm / hw, rem = mm - n * hw;
                    const int oy = rem / p.cWo, ox = rem - oy * p.cWo;
                    const int y0 = oy * p.cS - p.cP, x0 = ox * p.cS - p.cP;
                    poff[i] = (((long long)n * p.cH + y0) * p.cW + x0) * p.cC + 4 * j;
                    const int ylo = max(0, -y0), yhi = min(p.cF, p.cH - y0);
                    const int xlo = max(0, -x0), xhi = min(p.cF, p.cW - x0);
                    ymask[i] = (m < p.M && yhi > ylo) ? (((1u << yhi) - 1u) & ~((1u << ylo) - 1u)) : 0u;
                    xmask[i] = xhi > xlo ? (((1u << xhi) - 1u) & ~((1u << xlo) - 1u)) : 0u;
                }
            }
            int fy = 0, fx = 0, cch = 0;                                  // kConv: tap + channel chunk of kb
            const int nch = kConv ? p.cC / kBK : 1;
            for (int kb = 0; kb < p.nkb; ++kb) {
                mbar_wait(empty_bar + 8 * stage, phase ^ 1);
                const uint32_t a_stage = base + stage * kAStage;
                const int tap_off = kConv ? (fy * p.cW + fx) * p.cC + cch * kBK : 0;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int r = r0 + 16 * i, m = m0 + r;
                    const uint32_t sw = (uint32_t)(r & 7);
                    const uint32_t dst = a_stage + (uint32_t)(r >> 3) * 1024u + sw * 128u + (((uint32_t)j ^ sw) << 4);
                    if (kConv) {
                        const bool v = ((ymask[i] >> fy) & (xmask[i] >> fx) & 1u) != 0u;
                        cp_async16(dst, p.A + (v ? poff[i] + tap_off : 0), v ? 16u : 0u);
                    } else {
                        const float* src = p.A + (size_t)(m < p.M ? m : 0) * p.lda + kb * kBK + 4 * j;
                        cp_async16(dst, src, m < p.M ? 16u : 0u);
                    }
                }
                cp_async_mbar_arrive_noinc(full_bar + 8 * stage);
                if (kConv) {
                    if (++cch == nch) { cch = 0; if (++fx == p.cF) { fx = 0; ++fy; } }
                }
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 4) {
        // ------------------------------------------------ MMA issuer
        int stage = 0, buf = 0;
        uint32_t phase = 0, bphase = 0;
        for (long long t = blockIdx.x; t < total; t += gridDim.x) {
            mbar_wait(tempty_bar + 8 * buf, bphase ^ 1);
            tc_fence_after();
            for (int kb = 0; kb < p.nkb; ++kb) {
                mbar_wait(full_bar + 8 * stage, phase);
                tc_fence_after();
                if (lane == 0) {
                    const uint64_t da = make_smem_desc(base + stage * kAStage);
                    const uint64_t db = make_smem_desc(base + b_off + stage * kBStage);
#pragma unroll
                    for (int ks = 0; ks < kBK / 8; ++ks)
                        umma_tf32(tmem_base + (uint32_t)(buf * kBN), da + 2 * ks, db + 2 * ks, p.idesc,
                                  (kb | ks) != 0 ? 1u : 0u);
                    umma_commit(empty_bar + 8 * stage);
                    if (kb == p.nkb - 1) umma_commit(tfull_bar + 8 * buf);
                }
                __syncwarp();
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
            if (++buf == 2) { buf = 0; bphase ^= 1; }
        }
    } else if (warp == 5) {
        // ------------------------------------------------ TMA issuer (weights; rows >= N read as 0)
        int stage = 0;
        uint32_t phase = 0;
        for (long long t = blockIdx.x; t < total; t += gridDim.x) {
            const int n0 = (int)(t % p.tiles_n) * kBN;
            for (int kb = 0; kb < p.nkb; ++kb) {
                mbar_wait(empty_bar + 8 * stage, phase ^ 1);
                if (lane == 0) {
                    mbar_arrive_expect_tx(full_bar + 8 * stage, kBStage);
                    tma_load_2d(base + b_off + stage * kBStage, &tmapW, kb * kBK, n0, full_bar + 8 * stage);
                }
                __syncwarp();
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
        }
    } else {
        // ------------------------------------------------ epilogue (quarter = warp % 4)
        const int q = warp & 3;
        int buf = 0;
        uint32_t bphase = 0;
        for (long long t = blockIdx.x; t < total; t += gridDim.x) {
            const int m0 = (int)(t / p.tiles_n) * kBM, n0 = (int)(t % p.tiles_n) * kBN;
            const int m = m0 + q * 32 + lane;
            mbar_wait(tfull_bar + 8 * buf, bphase);
            tc_fence_after();
            for (int cc = 0; cc < kBN / 32; ++cc) {
                uint32_t r[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * kBN + cc * 32), r);
                const int nb = n0 + cc * 32;
                if (m < p.M && nb < p.N) {
                    float* crow = p.C + (size_t)m * p.ldc + nb;
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        if (nb + c >= p.N) break;
                        float v = __uint_as_float(r[c]);
                        if (p.flags & 8) v += p.bias[nb + c];
                        if (p.flags & 1) v = fmaxf(v, 0.f);
                        if (p.flags & 2) v = to_tf32(v);
                        if (p.flags & 4) atomicAdd(crow + c, v);          // RED.ADD.F32 (local or peer)
                        else crow[c] = v;
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar + 8 * buf);
            if (++buf == 2) { buf = 0; bphase ^= 1; }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 2 * kBN);
    }
}

}  // namespace

// ------------------------------------------------------------------------------- host ----
namespace {

struct TcGemm {
    int N, K, device;
    float* d_w;
    CUtensorMap tmap;
};

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                             const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                             CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn encode_fn() {
    static EncodeFn fn = nullptr;
    if (!fn) {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeFn)f;
    }
    return fn;
}

inline float round_tf32(float x) {
    uint32_t u;
    memcpy(&u, &x, 4);
    if ((u & 0x7F800000u) != 0x7F800000u) u = (u + 0x1000u) & ~0x1FFFu;
    memcpy(&x, &u, 4);
    return x;
}

}  // namespace

// w_nk: host weights [N][K] (K contiguous), i.e. torch's nn.Linear layout.  K must be a multiple of 32.
extern "C" int b200_tc_gemm_create(const float* w_nk, int N, int K, void** handle) {
    if (!w_nk || !handle || N <= 0 || K <= 0) return B200_ERR_ARG;
    if (K % kBK != 0) return B200_ERR_UNSUPPORTED;
    EncodeFn enc = encode_fn();
    if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return B200_ERR_CUDA; }
    TcGemm* t = new TcGemm();
    t->N = N; t->K = K;
    B200_CUDA_TRY(cudaGetDevice(&t->device));
    const int Npad = ((N + kBN - 1) / kBN) * kBN;             // whole TMA boxes: rows >= N are zeros
    std::vector<float> w((size_t)Npad * K, 0.f);
    for (size_t i = 0; i < (size_t)N * K; ++i) w[i] = round_tf32(w_nk[i]);
    B200_CUDA_TRY(cudaMalloc(&t->d_w, w.size() * sizeof(float)));
    B200_CUDA_TRY(cudaMemcpy(t->d_w, w.data(), w.size() * sizeof(float), cudaMemcpyHostToDevice));
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)Npad};
    cuuint64_t strides[1] = {(cuuint64_t)K * sizeof(float)};
    cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)kBN};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&t->tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, t->d_w, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed: %d (N=%d K=%d)", (int)r, N, K);
        cudaFree(t->d_w);
        delete t;
        return B200_ERR_CUDA;
    }
    *handle = t;
    return B200_OK;
}

extern "C" int b200_tc_gemm_destroy(void* handle) {
    TcGemm* t = (TcGemm*)handle;
    if (t) { cudaFree(t->d_w); delete t; }
    return B200_OK;
}

// flags: 1 ReLU, 2 round outputs to TF32, 4 accumulate into C with red.add (C may be peer memory and must
// be zeroed or hold the running sum), 8 add bias.  A (lda floats per row, 16 B aligned rows) and C may be
// peer-mapped pointers.
extern "C" int b200_tc_gemm(void* handle, const float* A, int lda, int M, const float* bias, float* C,
                            int ldc, int flags, void* stream) {
    TcGemm* t = (TcGemm*)handle;
    if (!t || !A || !C || M < 0 || lda < t->K || ldc < t->N || (lda & 3)) return B200_ERR_ARG;
    if ((flags & 8) && !bias) return B200_ERR_ARG;
    if (M == 0) return B200_OK;
    GemmParams p;
    memset(&p, 0, sizeof(p));
    p.A = A; p.C = C; p.bias = bias;
    p.M = M; p.N = t->N; p.K = t->K; p.lda = lda; p.ldc = ldc; p.flags = flags;
    p.tiles_m = (M + kBM - 1) / kBM;
    p.tiles_n = (t->N + kBN - 1) / kBN;
    p.nkb = t->K / kBK;
    p.idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kBN >> 3) << 17) | ((128u >> 4) << 24);
    const size_t smem = 1024 + (size_t)kStages * (kAStage + kBStage) + 16 * kStages + 32 + 64;
    static int sms = 0;
    if (!sms) B200_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, t->device));
    B200_CUDA_TRY(cudaFuncSetAttribute(tc_gemm_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long total = (long long)p.tiles_m * p.tiles_n;
    const unsigned grid = (unsigned)(total < sms ? total : sms);
    tc_gemm_kernel<false><<<grid, kGemmThreads, smem, (cudaStream_t)stream>>>(t->tmap, p);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

// Convolution as an implicit GEMM on the same kernel: x is NHWC [N][H][W][C] (C a multiple of 32, may be a
// peer pointer), the handle holds weights [K_out][F*F*C] in (fy, fx, c) order, out is [N*Ho*Wo][ldc].
// Same flags as b200_tc_gemm.  No im2col buffer is written or read.
extern "C" int b200_tc_conv_gemm(void* handle, const float* x, int N, int H, int W, int C, int F, int S, int P,
                                 const float* bias, float* out, int ldc, int flags, void* stream) {
    TcGemm* t = (TcGemm*)handle;
    if (!t || !x || !out || N < 0 || H <= 0 || W <= 0 || C <= 0 || F <= 0 || S <= 0 || P < 0) return B200_ERR_ARG;
    if (C % kBK != 0 || F > 16) return B200_ERR_UNSUPPORTED;
    if (t->K != F * F * C || ldc < t->N) return B200_ERR_ARG;
    if ((flags & 8) && !bias) return B200_ERR_ARG;
    const int Ho = (H - F + 2 * P) / S + 1, Wo = (W - F + 2 * P) / S + 1;
    if (Ho <= 0 || Wo <= 0) return B200_ERR_ARG;
    const long long M = (long long)N * Ho * Wo;
    if (M == 0) return B200_OK;
    if (M > 0x7fffffffLL - kBM) return B200_ERR_RANGE;
    GemmParams p;
    memset(&p, 0, sizeof(p));
    p.A = x; p.C = out; p.bias = bias;
    p.M = (int)M; p.N = t->N; p.K = t->K; p.lda = 0; p.ldc = ldc; p.flags = flags;
    p.tiles_m = (int)((M + kBM - 1) / kBM);
    p.tiles_n = (t->N + kBN - 1) / kBN;
    p.nkb = t->K / kBK;
    p.idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kBN >> 3) << 17) | ((128u >> 4) << 24);
    p.cH = H; p.cW = W; p.cC = C; p.cF = F; p.cS = S; p.cP = P; p.cHo = Ho; p.cWo = Wo;
    const size_t smem = 1024 + (size_t)kStages * (kAStage + kBStage) + 16 * kStages + 32 + 64;
    static int sms = 0;
    if (!sms) B200_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, t->device));
    B200_CUDA_TRY(cudaFuncSetAttribute(tc_gemm_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long total = (long long)p.tiles_m * p.tiles_n;
    const unsigned grid = (unsigned)(total < sms ? total : sms);
    tc_gemm_kernel<true><<<grid, kGemmThreads, smem, (cudaStream_t)stream>>>(t->tmap, p);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}
