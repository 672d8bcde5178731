// Exact-FP32 SIMT fused block: conv + bias + ReLU + 3x3/2 max-pool (+ LRN) in one kernel.
//
// This is the validation twin of the tcgen05 path (tc_block.cu) and the V3-equivalent kernel set:
// the reference runs conv / relu / pool / lrn as four 1-thread-per-output kernels with an HBM
// round trip between each (ref v3_cuda_only/src/layers_cuda.cu:20-139); here one CTA owns a
// TRxTC patch of conv outputs for every output channel, keeps the accumulators in registers
// (thread = output channel, acc[TR*TC]), pools them in registers and applies LRN through a few
// KB of shared memory, so nothing but the pooled/normalised result ever reaches memory.
//
// The NHWC trick that makes one loop serve both convs: for a fixed filter row fy the F*C values a
// pixel needs are CONTIGUOUS in the input row ((x*S .. x*S+F-1) x C), and so are the weights in
// [F][F][C][K] order.  The inner loop therefore walks a flat (fx,c) index 4 floats at a time:
// one broadcast LDS.128 of the input feeds 4 FFMA per thread.
//
// Input rows come from up to 4 row segments (local slab or NVLink-mapped peer slabs, see
// b200_row_seg), which is how the same kernel does the scatter pull and the conv2 halo read.
#include "common.cuh"

namespace {

using namespace b200;

struct TileShape { int tr, tc; };

__host__ __device__ inline int cdiv(int a, int b) { return (a + b - 1) / b; }

template <int TR, int TC>
__global__ void __launch_bounds__(256)
simt_block_kernel(const float* __restrict__ wt,    // [F][F*C][K]
                  const float* __restrict__ bias,  // [K]
                  float* __restrict__ out, const b200_block_geom g, int tiles_r, int tiles_c,
                  int kgroups) {
    constexpr int PR = (TR - 1) / 2, PC = (TC - 1) / 2;
    extern __shared__ __align__(16) float smem[];

    const int F = g.F, S = g.S, C = g.C, K = g.K;
    const int FC = F * C;
    const int IR = (TR - 1) * S + F;                  // input patch rows
    const int ICf = ((TC - 1) * S + F) * C;           // input patch row length in floats
    const int pitch = (ICf + 3) & ~3;

    int t = blockIdx.x;
    const int kg = t % kgroups; t /= kgroups;
    const int tc_i = t % tiles_c; t /= tiles_c;
    const int tr_i = t % tiles_r; t /= tiles_r;
    const int n = t;

    const int pr0 = g.p0 + tr_i * PR, pc0 = tc_i * PC;        // first pooled row / col of the tile
    const int ir0 = pr0 * 2 * S - g.P;                         // input row / col of the patch origin
    const int ic0 = pc0 * 2 * S - g.P;

    // ---- wait for remote rows this patch touches (consumer side of the flag protocol) ----
    if (threadIdx.x == 0) {
        block_gate_wait(g);
        for (int s = 0; s < g.nseg; ++s) {
            const b200_row_seg& sg = g.seg[s];
            if (sg.flag && ir0 < sg.hi && ir0 + IR > sg.lo)
                wait_flag_ge(sg.flag + n, sg.target * (g.epoch ? *g.epoch : 1), g.status, g.timeout_ns);
        }
    }
    __syncthreads();

    // ---- stage the input patch (zero outside the image = conv zero padding) ----
    const bool vec_ok = (C % 4 == 0) && ((g.W * C) % 4 == 0);
    for (int r = 0; r < IR; ++r) {
        const int h = ir0 + r;
        const float* src = nullptr;
        if (h >= 0 && h < g.H) {
            for (int s = 0; s < g.nseg; ++s) {
                const b200_row_seg& sg = g.seg[s];
                if (h >= sg.lo && h < sg.hi) {
                    src = sg.ptr + ((size_t)n * sg.slab + (h - sg.base)) * (g.in_pitch ? g.in_pitch : g.W * C);
                    break;
                }
            }
        }
        float* dst = smem + r * pitch;
        if (vec_ok) {
            for (int q = threadIdx.x; q < pitch / 4; q += blockDim.x) {
                const int x = ic0 + (q * 4) / C;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (src && x >= 0 && x < g.W && q * 4 < ICf)
                    v = *reinterpret_cast<const float4*>(src + (size_t)x * C + (q * 4) % C);
                *reinterpret_cast<float4*>(dst + q * 4) = v;
            }
        } else {
            for (int f = threadIdx.x; f < pitch; f += blockDim.x) {
                const int x = ic0 + f / C;
                float v = 0.f;
                if (src && x >= 0 && x < g.W && f < ICf) v = src[(size_t)x * C + f % C];
                dst[f] = v;
            }
        }
    }
    __syncthreads();

    // ---- main loop: thread = output channel ----
    const int kpg = cdiv(K, kgroups);                         // channels per group
    const int k = kg * kpg + threadIdx.x;
    const bool kvalid = threadIdx.x < kpg && k < K;
    const int kc = kvalid ? k : 0;

    float acc[TR][TC];
#pragma unroll
    for (int i = 0; i < TR; ++i)
#pragma unroll
        for (int j = 0; j < TC; ++j) acc[i][j] = 0.f;

    const int SC = S * C;
    const int FCv = (SC % 4 == 0) ? (FC & ~3) : 0;            // LDS.128 needs 16 B aligned taps
    for (int fy = 0; fy < F; ++fy) {
        const float* wrow = wt + (size_t)fy * FC * K + kc;
        const float* prow = smem + fy * pitch;
        int tt = 0;
        for (; tt < FCv; tt += 4) {
            const float w0 = __ldg(wrow + (size_t)(tt + 0) * K);
            const float w1 = __ldg(wrow + (size_t)(tt + 1) * K);
            const float w2 = __ldg(wrow + (size_t)(tt + 2) * K);
            const float w3 = __ldg(wrow + (size_t)(tt + 3) * K);
#pragma unroll
            for (int i = 0; i < TR; ++i)
#pragma unroll
                for (int j = 0; j < TC; ++j) {
                    const float4 v = *reinterpret_cast<const float4*>(
                        prow + i * S * pitch + j * SC + tt);
                    acc[i][j] = fmaf(v.x, w0, acc[i][j]);
                    acc[i][j] = fmaf(v.y, w1, acc[i][j]);
                    acc[i][j] = fmaf(v.z, w2, acc[i][j]);
                    acc[i][j] = fmaf(v.w, w3, acc[i][j]);
                }
        }
        for (; tt < FC; ++tt) {
            const float w0 = __ldg(wrow + (size_t)tt * K);
#pragma unroll
            for (int i = 0; i < TR; ++i)
#pragma unroll
                for (int j = 0; j < TC; ++j)
                    acc[i][j] = fmaf(prow[i * S * pitch + j * SC + tt], w0, acc[i][j]);
        }
    }

    // ---- epilogue: bias + ReLU + pool in registers ----
    const float b = kvalid ? bias[k] : 0.f;
    float pooled[PR][PC];
#pragma unroll
    for (int pi = 0; pi < PR; ++pi)
#pragma unroll
        for (int pj = 0; pj < PC; ++pj) {
            float m = -INFINITY;
#pragma unroll
            for (int a = 0; a < 3; ++a)
#pragma unroll
                for (int c = 0; c < 3; ++c) m = fmaxf(m, acc[2 * pi + a][2 * pj + c]);
            pooled[pi][pj] = fmaxf(m + b, 0.f);        // relu(max(x)+b) == max(relu(x+b))
        }

    const bool do_lrn = g.lrn.N_lrn > 0;
    float* lbuf = smem;                                        // reuse: [PR*PC][K] pooled values
    if (do_lrn) {
        __syncthreads();                                       // everyone is done with the patch
        if (kvalid) {
#pragma unroll
            for (int pi = 0; pi < PR; ++pi)
#pragma unroll
                for (int pj = 0; pj < PC; ++pj) lbuf[(pi * PC + pj) * K + k] = pooled[pi][pj];
        }
        __syncthreads();
    }
    if (kvalid) {
        const int half = g.lrn.N_lrn / 2;
        const float a = g.lrn.mode == B200_LRN_ALPHA_OVER_N ? g.lrn.alpha / (float)g.lrn.N_lrn
                                                            : g.lrn.alpha;
#pragma unroll
        for (int pi = 0; pi < PR; ++pi) {
            const int pr = pr0 + pi;
            if (pr >= g.p1 || pr >= g.Hp) continue;
#pragma unroll
            for (int pj = 0; pj < PC; ++pj) {
                const int pc = pc0 + pj;
                if (pc >= g.Wp) continue;
                float v = pooled[pi][pj];
                if (do_lrn) {
                    float ss = 0.f;
                    const int c0 = max(0, k - half), c1 = min(K - 1, k + half);
                    for (int c = c0; c <= c1; ++c) {
                        const float x = lbuf[(pi * PC + pj) * K + c];
                        ss = fmaf(x, x, ss);
                    }
                    v = v / powf(g.lrn.k + a * ss, g.lrn.beta);
                }
                out[(((size_t)n * g.out_slab + (pr - g.out_base)) * g.Wp + pc) * K + k] = v;
                for (int m = 0; m < g.n_mirror; ++m)           // halo push: the neighbour's copy of this row
                    if (pr >= g.mirror[m].lo && pr < g.mirror[m].hi)
                        g.mirror[m].ptr[(((size_t)n * g.mirror[m].slab + (pr - g.mirror[m].base)) * g.Wp + pc) * K + k] = v;
            }
        }
    }

    // ---- producer side of the flag protocol ----
    if (g.n_signal > 0 || g.done_count) {
        __syncthreads();
        if (threadIdx.x == 0) {
            tile_signal(g, n, pr0, min(pr0 + PR, g.p1));
            block_done_signal(g);
        }
    }
}

// Tile shape per layer: 7x7 conv patch (3x3 pooled) by default; a 5x15 patch (2x7 pooled) when the
// pooled width is not a multiple of 3 but fits 7-wide bands better (13 = 7+6).
TileShape pick_tile(const b200_block_geom* g) {
    const int w3 = cdiv(g->Wp, 3) * 3, w7 = cdiv(g->Wp, 7) * 7;
    // cost ~ conv pixels computed per pooled row band
    const double c3 = (double)w3 / 3 * 49 / 3.0, c7 = (double)w7 / 7 * 75 / 2.0;
    return c7 < c3 ? TileShape{5, 15} : TileShape{7, 7};
}

int kgroups_for(int K) { return cdiv(K, 256); }

}  // namespace

extern "C" int b200_simt_block_tiles(const b200_block_geom* g) {
    if (!g) return B200_ERR_ARG;
    TileShape ts = pick_tile(g);
    const int PR = (ts.tr - 1) / 2, PC = (ts.tc - 1) / 2;
    const int rows = g->p1 - g->p0;
    if (rows <= 0) return 0;
    return cdiv(rows, PR) * cdiv(g->Wp, PC) * kgroups_for(g->K);
}

extern "C" int b200_simt_block_tiles_rows(const b200_block_geom* g, int lo, int hi) {
    if (!g) return B200_ERR_ARG;
    TileShape ts = pick_tile(g);
    const int PR = (ts.tr - 1) / 2, PC = (ts.tc - 1) / 2;
    int n = 0;
    for (int r0 = g->p0; r0 < g->p1; r0 += PR) {
        const int r1 = r0 + PR < g->p1 ? r0 + PR : g->p1;
        if (hi <= lo || (r0 < hi && r1 > lo)) ++n;
    }
    return n * cdiv(g->Wp, PC) * kgroups_for(g->K);
}

extern "C" int b200_simt_block(const float* wt, const float* bias, float* out,
                               const b200_block_geom* g, void* stream) {
    if (!g || !wt || !bias || !out) return B200_ERR_ARG;
    if (g->nseg < 1 || g->nseg > B200_MAX_SEG || g->n_signal < 0 || g->n_signal > B200_MAX_SIGNAL ||
        g->n_done < 0 || g->n_done > 2 || g->n_mirror < 0 || g->n_mirror > B200_MAX_MIRROR)
        return B200_ERR_ARG;
    const int rows = g->p1 - g->p0;
    if (rows <= 0 || g->N <= 0) return B200_OK;
    if (g->lrn.N_lrn > 0 && g->K > 256) return B200_ERR_UNSUPPORTED;   // LRN needs all K in one CTA
    TileShape ts = pick_tile(g);
    const int PR = (ts.tr - 1) / 2, PC = (ts.tc - 1) / 2;
    const int tiles_r = cdiv(rows, PR), tiles_c = cdiv(g->Wp, PC);
    const int kgroups = kgroups_for(g->K);
    const int kpg = cdiv(g->K, kgroups);
    const int threads = cdiv(kpg, 32) * 32;
    const int IR = (ts.tr - 1) * g->S + g->F;
    const int pitch = ((((ts.tc - 1) * g->S + g->F) * g->C) + 3) & ~3;
    size_t smem = (size_t)IR * pitch * sizeof(float);
    const size_t lrn_smem = (size_t)PR * PC * g->K * sizeof(float);
    if (lrn_smem > smem) smem = lrn_smem;
    if (smem > 227 * 1024) return B200_ERR_UNSUPPORTED;
    dim3 grid((unsigned)((size_t)g->N * tiles_r * tiles_c * kgroups));
    cudaStream_t st = (cudaStream_t)stream;
    if (ts.tr == 7) {
        B200_CUDA_TRY(cudaFuncSetAttribute(simt_block_kernel<7, 7>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        simt_block_kernel<7, 7><<<grid, threads, smem, st>>>(wt, bias, out, *g, tiles_r, tiles_c, kgroups);
    } else {
        B200_CUDA_TRY(cudaFuncSetAttribute(simt_block_kernel<5, 15>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        simt_block_kernel<5, 15><<<grid, threads, smem, st>>>(wt, bias, out, *g, tiles_r, tiles_c, kgroups);
    }
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}
