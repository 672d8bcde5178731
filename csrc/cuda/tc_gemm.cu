// tcgen05 GEMM for the full-AlexNet extension: C[M][N] (+)= act(A[M][K] * W[N][K]^T + bias).
//
// The reference stops after LRN2; FC6-8 (and conv3-5) are only planned there
// (ref final_project/PROBLEMS.txt:188-190, summary.md:37-43,68).  This kernel is the building block
// of that extension and of its tensor-parallel form:
//   * A is gathered by 4 producer warps with cp.async straight into the 128B-swizzled stage.  A is a
//     plain pointer, so it may be another GPU's activation buffer mapped over NVLink: the
//     all-gather of "all-gather -> GEMM" is the loader itself.
//   * W (K-major, TF32-rounded once) arrives by TMA; tcgen05.mma kind::tf32 M128 x N x K8 with N = 256 when
//     the layer has more than 128 outputs (a tcgen05.mma costs ~160 cycles whatever N is:
//     profiles/ncu/final_r1_summary.txt), else 128; fp32 accumulators double-buffered in TMEM (2 x N
//     columns) so tile t's epilogue overlaps tile t+1's main loop.
//   * Small-M layers (FC at batch <= 256 fill a fraction of the 148 SMs) can be split along K: split ks of
//     a tile writes its partial sums to C + ks * split_stride and b200_splitk_reduce (ext_ops.cu) adds
//     them in a fixed order (deterministic, unlike an atomic split-K).
//   * Epilogue: tcgen05.ld -> +bias -> ReLU -> either plain stores, or red.global.add.f32 into C, which
//     may be a peer pointer: the reduce(-scatter) of "GEMM -> reduce-scatter" is the epilogue itself.
// Warps: 0-3 A producers, 4 MMA issuer + TMEM owner, 5 TMA issuer, 6-9 epilogue (TMEM quarter = warp%4).
#include "tc_ptx.cuh"

#include <cstdlib>
#include <cstring>
#include <vector>

namespace {

using namespace b200;
using namespace b200::tc;

constexpr int kBM = 128, kBK = 32;
constexpr int kAStage = kBM * 128;                               // bytes
constexpr int kGemmThreads = 320;
__host__ __device__ constexpr int gemm_stages(int bn) { return bn == 256 ? 4 : 5; }  // 4 x 48 KB or 5 x 32 KB
constexpr size_t gemm_smem(int bn) {
    return 1024 + (size_t)gemm_stages(bn) * (kAStage + bn * 128) + 16 * gemm_stages(bn) + 32 + 64;
}
constexpr int kProdThreads = 128;

struct GemmParams {
    const float* A;
    float* C;
    const float* bias;
    int M, N, K, lda, ldc;
    int flags;                                                   // 1 relu, 2 round to tf32, 4 atomic add, 8 bias
    int tiles_m, tiles_n, nkb;
    int ksplit, kb_per;                                          // split-K: k-blocks [ks*kb_per, ..) per split
    long long split_stride;                                      // floats between the partial C of two splits
    uint32_t idesc;
    // implicit-GEMM convolution (kConv): A is an NHWC activation [cN][cH][cW][cC], row m of the GEMM is
    // output pixel m = (n*cHo + oy)*cWo + ox and k-block kb is 32 channels of filter tap (fy, fx):
    // kb = (fy*cF + fx)*(cC/32) + chunk.  Taps outside the image are zero-filled by cp.async (size 0).
    int cH, cW, cC, cF, cS, cP, cHo, cWo;
};

// kConv = false: A[M][lda] row-major.  kConv = true: A rows are gathered from the NHWC activation, so the
// im2col matrix of conv3-5 never exists in memory.
struct GemmTile { int m0, n0, kb0, kb1, ks; };

// tile index -> (m tile, k split, n tile), n fastest so that concurrently running CTAs share A rows in L2
__device__ __forceinline__ GemmTile gemm_tile(const GemmParams& p, long long t, int bn) {
    GemmTile g;
    const int tn = (int)(t % p.tiles_n);
    const long long u = t / p.tiles_n;
    g.ks = (int)(u % p.ksplit);
    g.m0 = (int)(u / p.ksplit) * kBM;
    g.n0 = tn * bn;
    g.kb0 = g.ks * p.kb_per;
    g.kb1 = min(p.nkb, g.kb0 + p.kb_per);
    return g;
}

template <bool kConv, int kBN>
__global__ void __launch_bounds__(kGemmThreads, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tmapW, const __grid_constant__ GemmParams p) {
    constexpr int kStages = gemm_stages(kBN);
    constexpr int kBStage = kBN * 128;
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
    const long long total = (long long)p.tiles_m * p.tiles_n * p.ksplit;

    if (warp < 4) {
        // ------------------------------------------------ A producers (local or peer memory)
        const int tid = threadIdx.x, j = tid & 7, r0 = tid >> 3;          // rows r0 + 16*i
        int stage = 0;
        uint32_t phase = 0;
        for (long long t = blockIdx.x; t < total; t += gridDim.x) {
            const GemmTile gt = gemm_tile(p, t, kBN);
            const int m0 = gt.m0;
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
            const int nch = kConv ? p.cC / kBK : 1;
            int cch = gt.kb0 % nch;                                       // kConv: channel chunk + tap of kb
            int fx = (gt.kb0 / nch) % (kConv ? p.cF : 1), fy = kConv ? (gt.kb0 / nch) / p.cF : 0;
            for (int kb = gt.kb0; kb < gt.kb1; ++kb) {
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
            const GemmTile gt = gemm_tile(p, t, kBN);
            mbar_wait(tempty_bar + 8 * buf, bphase ^ 1);
            tc_fence_after();
            for (int kb = gt.kb0; kb < gt.kb1; ++kb) {
                mbar_wait(full_bar + 8 * stage, phase);
                fence_proxy_async_smem();                       // producers' cp.async data -> async-proxy reads
                tc_fence_after();
                if (lane == 0) {
                    const uint64_t da = make_smem_desc(base + stage * kAStage);
                    const uint64_t db = make_smem_desc(base + b_off + stage * kBStage);
#pragma unroll
                    for (int ks = 0; ks < kBK / 8; ++ks)
                        umma_tf32(tmem_base + (uint32_t)(buf * kBN), da + 2 * ks, db + 2 * ks, p.idesc,
                                  (kb != gt.kb0 || ks != 0) ? 1u : 0u);
                    umma_commit(empty_bar + 8 * stage);
                    if (kb == gt.kb1 - 1) umma_commit(tfull_bar + 8 * buf);
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
            const GemmTile gt = gemm_tile(p, t, kBN);
            const int n0 = gt.n0;
            for (int kb = gt.kb0; kb < gt.kb1; ++kb) {
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
        const bool vec_ok = !(p.flags & 4) && (p.ldc & 3) == 0 && (p.split_stride & 3) == 0 &&
                            (reinterpret_cast<uintptr_t>(p.C) & 15) == 0;
        for (long long t = blockIdx.x; t < total; t += gridDim.x) {
            const GemmTile gt = gemm_tile(p, t, kBN);
            const int m0 = gt.m0, n0 = gt.n0;
            const int m = m0 + q * 32 + lane;
            float* cbase = p.C + (long long)gt.ks * p.split_stride;
            mbar_wait(tfull_bar + 8 * buf, bphase);
            tc_fence_after();
            for (int cc = 0; cc < kBN / 32; ++cc) {
                const int nb = n0 + cc * 32;
                if (nb >= p.N) break;                                     // warp-uniform: rest of the tile is padding
                uint32_t r[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * kBN + cc * 32), r);
                if (m < p.M) {
                    float* crow = cbase + (size_t)m * p.ldc + nb;
                    if (vec_ok && nb + 32 <= p.N) {                       // whole chunk: 8 x 16 B stores per row
#pragma unroll
                        for (int c = 0; c < 32; c += 4) {
                            float v[4];
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                v[e] = __uint_as_float(r[c + e]);
                                if (p.flags & 8) v[e] += p.bias[nb + c + e];
                                if (p.flags & 1) v[e] = fmaxf(v[e], 0.f);
                                if (p.flags & 2) v[e] = to_tf32(v[e]);
                            }
                            *reinterpret_cast<float4*>(crow + c) = make_float4(v[0], v[1], v[2], v[3]);
                        }
                    } else {
#pragma unroll
                        for (int c = 0; c < 32; ++c) {
                            if (nb + c >= p.N) break;
                            float v = __uint_as_float(r[c]);
                            if (p.flags & 8) v += p.bias[nb + c];
                            if (p.flags & 1) v = fmaxf(v, 0.f);
                            if (p.flags & 2) v = to_tf32(v);
                            if (p.flags & 4) atomicAdd_system(crow + c, v);   // RED.ADD.F32 at system scope: C may be a peer's
                            else crow[c] = v;
                        }
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
    int N, K, device, sms;
    float* d_w;
    CUtensorMap tmap128, tmap256;                      // TMA boxes of 128 / 256 weight rows
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

// N tile: 256 when the layer is wider than one 128 tile (B200_GEMM_BN128=1 forces the 128 kernel)
inline int pick_bn(int N) {
    static const bool bn128 = getenv("B200_GEMM_BN128") != nullptr;     // read once, not on the launch path
    return (N > 128 && !bn128) ? 256 : 128;
}

template <bool kConv, int kBN>
int launch_one(const TcGemm* t, const GemmParams& p, unsigned grid, cudaStream_t stream) {
    const size_t smem = gemm_smem(kBN);
    B200_CUDA_TRY(cudaFuncSetAttribute(tc_gemm_kernel<kConv, kBN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    tc_gemm_kernel<kConv, kBN><<<grid, kGemmThreads, smem, stream>>>(kBN == 256 ? t->tmap256 : t->tmap128, p);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

// fills the tiling fields of p (M, N, K, flags, pointers already set) and launches
int launch_gemm(const TcGemm* t, GemmParams& p, bool conv, int ksplit, long long split_stride, void* stream) {
    const int bn = pick_bn(t->N);
    p.tiles_m = (p.M + kBM - 1) / kBM;
    p.tiles_n = (t->N + bn - 1) / bn;
    p.nkb = t->K / kBK;
    if (ksplit < 1) ksplit = 1;
    if (ksplit > p.nkb) ksplit = p.nkb;
    p.kb_per = (p.nkb + ksplit - 1) / ksplit;
    p.ksplit = (p.nkb + p.kb_per - 1) / p.kb_per;      // every split owns at least one k-block
    p.split_stride = p.ksplit > 1 ? split_stride : 0;
    p.idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(bn >> 3) << 17) | ((128u >> 4) << 24);
    const int sms = t->sms;
    const long long total = (long long)p.tiles_m * p.tiles_n * p.ksplit;
    const unsigned grid = (unsigned)(total < sms ? total : sms);
    cudaStream_t st = (cudaStream_t)stream;
    if (bn == 256) return conv ? launch_one<true, 256>(t, p, grid, st) : launch_one<false, 256>(t, p, grid, st);
    return conv ? launch_one<true, 128>(t, p, grid, st) : launch_one<false, 128>(t, p, grid, st);
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
    B200_CUDA_TRY(cudaDeviceGetAttribute(&t->sms, cudaDevAttrMultiProcessorCount, t->device));
    const int Npad = ((N + 255) / 256) * 256;                 // whole TMA boxes: rows >= N are zeros
    std::vector<float> w((size_t)Npad * K, 0.f);
    for (size_t i = 0; i < (size_t)N * K; ++i) w[i] = round_tf32(w_nk[i]);
    B200_CUDA_TRY(cudaMalloc(&t->d_w, w.size() * sizeof(float)));
    B200_CUDA_TRY(cudaMemcpy(t->d_w, w.data(), w.size() * sizeof(float), cudaMemcpyHostToDevice));
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)Npad};
    cuuint64_t strides[1] = {(cuuint64_t)K * sizeof(float)};
    cuuint32_t estr[2] = {1, 1};
    for (int rows = 128; rows <= 256; rows *= 2) {
        cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)rows};
        CUresult r = enc(rows == 128 ? &t->tmap128 : &t->tmap256, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, t->d_w, dims,
                         strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            set_error("cuTensorMapEncodeTiled failed: %d (N=%d K=%d box rows %d)", (int)r, N, K, rows);
            cudaFree(t->d_w);
            delete t;
            return B200_ERR_CUDA;
        }
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
// peer-mapped pointers.  ksplit > 1: split ks of every tile covers a slice of K and stores its partial sums
// at C + ks * split_stride (flags must be 0; finish with b200_splitk_reduce, which applies bias/activation).
extern "C" int b200_tc_gemm_splitk(void* handle, const float* A, int lda, int M, const float* bias, float* C,
                                   int ldc, int flags, int ksplit, long long split_stride, void* stream) {
    TcGemm* t = (TcGemm*)handle;
    if (!t || !A || !C || M < 0 || lda < t->K || ldc < t->N || (lda & 3)) return B200_ERR_ARG;
    if ((flags & 8) && !bias) return B200_ERR_ARG;
    if (ksplit > 1 && (flags != 0 || split_stride < (long long)M * ldc)) return B200_ERR_ARG;
    if (ksplit > 1) {       // the caller sizes its workspace and its reduce by ksplit: never run another count silently
        const int nkb = t->K / kBK, kb_per = (nkb + ksplit - 1) / ksplit;
        if (ksplit > nkb || (nkb + kb_per - 1) / kb_per != ksplit) {
            set_error("tc_gemm_splitk: %d k-blocks cannot be cut into exactly %d splits of %d (that gives %d); "
                      "ask for the effective count", nkb, ksplit, kb_per, (nkb + kb_per - 1) / kb_per);
            return B200_ERR_ARG;
        }
    }
    if (M == 0) return B200_OK;
    GemmParams p;
    memset(&p, 0, sizeof(p));
    p.A = A; p.C = C; p.bias = bias;
    p.M = M; p.N = t->N; p.K = t->K; p.lda = lda; p.ldc = ldc; p.flags = flags;
    return launch_gemm(t, p, false, ksplit, split_stride, stream);
}

extern "C" int b200_tc_gemm(void* handle, const float* A, int lda, int M, const float* bias, float* C,
                            int ldc, int flags, void* stream) {
    return b200_tc_gemm_splitk(handle, A, lda, M, bias, C, ldc, flags, 1, 0, stream);
}

// how many k-blocks (of 32) the handle's K has, and the N tile the kernel will use: lets the caller size a
// split-K workspace and pick ksplit from the tile count
extern "C" int b200_tc_gemm_info(void* handle, int* nkb, int* bn) {
    TcGemm* t = (TcGemm*)handle;
    if (!t) return B200_ERR_ARG;
    if (nkb) *nkb = t->K / kBK;
    if (bn) *bn = pick_bn(t->N);
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
    p.cH = H; p.cW = W; p.cC = C; p.cF = F; p.cS = S; p.cP = P; p.cHo = Ho; p.cWo = Wo;
    return launch_gemm(t, p, true, 1, 0, stream);
}
