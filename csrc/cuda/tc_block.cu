// tcgen05 / TMEM / TMA fused conv block for sm_100a:
//   implicit-GEMM conv (kind::tf32, fp32 accumulate in TMEM) + bias + ReLU + 3x3/2 max-pool
//   [+ cross-channel LRN], with the scatter pull / halo read / gather store over peer memory.
//
// Replaces, per block, the reference's convKernel + reluKernel + poolKernel (+ lrnKernel)
// (ref v3_cuda_only/src/layers_cuda.cu:20-139, v4_mpi_cuda/src/layers_mpi_cuda.cu:25-89) and the
// host-staged MPI traffic around them (ref v4_mpi_cuda/src/main_mpi_cuda.cpp:58-79,100-130).
//
// GEMM view.  M = conv output pixels, N = output channels (all of them in one CTA so that LRN,
// a window over channels, can run in the epilogue), K = F*F*C ordered (fy, fx, c).
//   * A (im2col rows) never exists in memory.  8 producer warps gather it straight from the NHWC
//     activation — local slab or NVLink-mapped peer slab, chosen per input row through
//     b200_row_seg — round to TF32 and store it into shared memory in the canonical K-major
//     128B-swizzled UMMA layout, then fence.proxy.async + mbarrier arrive.
//   * B (weights, K-major [N][Kpad]) arrives by TMA (cp.async.bulk.tensor.2d, SWIZZLE_128B) on the
//     same mbarrier (expect_tx).
//   * One elected thread issues tcgen05.mma.cta_group::1.kind::tf32 (M=128 x N x K=8); a tile is
//     up to 256 pixels = two M=128 accumulators sharing every B stage (halves B traffic per FLOP).
//     tcgen05.commit releases the stage / signals the epilogue.
//   * A tile is a TR x TC patch of conv pixels = (2PR+1) x (2PC+1), so PR x PC pooled outputs are
//     complete inside the tile.  Epilogue: tcgen05.ld 32 columns at a time -> smem patch ->
//     max-pool -> (+bias, ReLU; they commute with max) -> LRN over the channel axis -> store.
//     The store target is whatever `out` points at: for block 2 that is rank 0's output tensor
//     (the gather); the epilogue also bumps arrival counters on peers that will read this tile.
//
// Warp roles (320 threads): warps 0-7 producers, then epilogue; warp 8 MMA issuer + TMEM owner;
// warp 9 TMA issuer.  Persistent: grid = min(tiles, SMs), static round-robin over tiles.
#include "tc_ptx.cuh"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace {

using namespace b200;

constexpr int kProducerWarps = 8;
constexpr int kProducerThreads = kProducerWarps * 32;
constexpr int kMmaWarp = 8;
constexpr int kTmaWarp = 9;
constexpr int kThreads = 320;
constexpr int kBlockK = 32;                        // fp32 per k-block = one 128 B swizzle row
constexpr int kTileM = 256;                        // pixel slots per tile (2 x UMMA_M)
constexpr int kAStageBytes = kTileM * 128;         // 32 KB
constexpr int kCbufPitch = 33;                     // floats; conflict-free row/column access
constexpr int kCbufBytes = kTileM * kCbufPitch * 4;
constexpr int kMaxPooled = 64;                     // pooled outputs per tile

constexpr int kRowTab = 192;                       // row-pointer table entries (vec mode)

struct TcParams {
    b200_block_geom g;
    int TR, TC, PR, PC;
    int tiles_r, tiles_c;
    int nacc;
    int nkb;
    int stages, b_stage_bytes;
    int tmem_cols;
    int any_flag;
    int dbg;                     // perf experiments only: 1 skip A copies, 2 skip B TMA, 4 skip epilogue math;
                                 // conv1 kernel: 64 skip pixel TMA, 128 skip weight TMA, 256 skip epilogue, 512 skip MMAs, 4096 skip the tail copies;
                                 // plan-time: 1024 keep the 3-stage ring, 2048 pool through smem strips
    int round_out;               // round stored outputs to TF32 (next layer's A operand is then exact)
    int FCpad;                   // patch mode: floats per filter row in K (F*C rounded up to 4)
    int patch_rows, patch_pitch; // patch mode: staged input patch (rows incl. 1 guard row, floats per row)
    uint32_t patch_off, ctl_off;
    uint32_t idesc;
    long long total_tiles;
};

using namespace b200::tc;

__device__ __forceinline__ const float* seg_row_ptr(const b200_block_geom& g, int n, int h) {
#pragma unroll
    for (int s = 0; s < B200_MAX_SEG; ++s) {
        if (s < g.nseg && h >= g.seg[s].lo && h < g.seg[s].hi)
            return g.seg[s].ptr + ((size_t)n * g.seg[s].slab + (h - g.seg[s].base)) * (g.in_pitch ? g.in_pitch : g.W * g.C);
    }
    return nullptr;
}

// ------------------------------------------------------ shared fused epilogue ----
// TMEM -> shared memory -> 3x3/2 max-pool -> +bias, ReLU [-> LRN over channels] -> store (+ halo push), used by
// every kernel whose accumulator is pixels x channels (tc_block_kernel, tc_block_db_kernel, tc_block_pair_kernel).
// kWarps epilogue warps (4: one 128-pixel accumulator; 8: two) synchronise on named barrier kBar; warp `we`
// drains TMEM lane quarter we % 4 of accumulator we / 4, 32 channels at a time, into a double-buffered chunk
// buffer; each warp then pools a subset of the pooled pixels (lane = channel inside the chunk).
// `release()` runs once every tcgen05.ld of the calling warp has completed — the double-buffered kernels hand
// the accumulator back to the MMA warp there, BEFORE the LRN pass.
__device__ __forceinline__ float* out_ptr_at(float* base, int n, int slab, int row0, int pr, int Wp, int pc, int N) {
    return base + (((size_t)n * slab + (pr - row0)) * Wp + pc) * N;
}

template <int kBar, int kWarps>
__device__ __forceinline__ void bar_group() {
    asm volatile("bar.sync %0, %1;" ::"n"(kBar), "n"(kWarps * 32) : "memory");
}

template <int kBar, int kWarps, typename Release>
__device__ __forceinline__ void epilogue_pool_lrn(const TcParams& p, uint32_t tmem_acc, float* cbuf0, int cbuf_floats,
                                                  float* pbuf, const int2* ptab, const float* s_bias, float* out,
                                                  int n, int pr0, int pc0, int we, int lane, bool has_acc,
                                                  Release&& release) {
    const b200_block_geom& g = p.g;
    const int N = g.K;
    const int q = we & 3, hacc = we >> 2;
    const int te = we * 32 + lane;
    const int row = hacc * 128 + q * 32 + lane;              // pixel slot this thread drains
    const int PP = p.PR * p.PC, nchunk = N / 32, NP = N + 4; // pbuf row: [2 zeros | N channels | 2 zeros]
    const bool do_lrn = g.lrn.N_lrn > 0;
    for (int cc = 0; cc < ((p.dbg & 4) ? 0 : nchunk); ++cc) {
        float* cb = cbuf0 + (cc & 1) * cbuf_floats;
        if (has_acc) {
            uint32_t r[32];
            tmem_ld32(tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)(hacc * N + cc * 32), r);
#pragma unroll
            for (int c = 0; c < 32; ++c) cb[row * kCbufPitch + c] = __uint_as_float(r[c]);
        }
        bar_group<kBar, kWarps>();                            // chunk cc is whole; chunk cc-1's reads are over
        const float bch = s_bias[cc * 32 + lane];
        for (int pp = we; pp < ((p.dbg & 32) ? 0 : PP); pp += kWarps) {
            const int2 e = ptab[pp];
            const float* c0p = cb + e.x + lane;
            float mx = -INFINITY;
#pragma unroll
            for (int a = 0; a < 3; ++a)
#pragma unroll
                for (int b = 0; b < 3; ++b) mx = fmaxf(mx, c0p[(a * p.TC + b) * kCbufPitch]);
            float val = fmaxf(mx + bch, 0.f);                 // relu(max(x)+b) == max(relu(x+b))
            if (do_lrn) {
                pbuf[pp * NP + 2 + cc * 32 + lane] = val;
            } else {
                const int pr = pr0 + (e.y >> 16), pc = pc0 + (e.y & 0xFFFF);
                if (p.round_out) val = to_tf32(val);
                if (pr < g.p1 && pr < g.Hp && pc < g.Wp) {
                    out_ptr_at(out, n, g.out_slab, g.out_base, pr, g.Wp, pc, N)[cc * 32 + lane] = val;
                    for (int m = 0; m < g.n_mirror; ++m)      // halo push: the neighbour's copy of this row
                        if (pr >= g.mirror[m].lo && pr < g.mirror[m].hi)
                            out_ptr_at(g.mirror[m].ptr, n, g.mirror[m].slab, g.mirror[m].base, pr, g.Wp, pc,
                                       N)[cc * 32 + lane] = val;
                }
            }
        }
    }
    tc_fence_before();
    __syncwarp();
    release();
    if (do_lrn && !(p.dbg & (4 | 16))) {
        bar_group<kBar, kWarps>();                            // pbuf is whole
        const int half = g.lrn.N_lrn / 2;
        const float al = g.lrn.mode == B200_LRN_ALPHA_OVER_N ? g.lrn.alpha / (float)max(g.lrn.N_lrn, 1) : g.lrn.alpha;
        const float nbeta = -g.lrn.beta;
        const bool win5 = g.lrn.N_lrn == 5;                   // AlexNet's window: straight-line 5 taps
        for (int pp0 = 0; pp0 < PP; pp0 += 4) {               // 4 independent pixels per trip (ILP)
            float* optr[4];
            const float* prow[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int pp = pp0 + u < PP ? pp0 + u : PP - 1;
                const int2 e = ptab[pp];
                const int pr = pr0 + (e.y >> 16), pc = pc0 + (e.y & 0xFFFF);
                const bool ok = pp0 + u < PP && pr < g.p1 && pr < g.Hp && pc < g.Wp;
                optr[u] = ok ? out_ptr_at(out, n, g.out_slab, g.out_base, pr, g.Wp, pc, N) : nullptr;
                prow[u] = pbuf + pp * NP;                     // channel c sits at prow[c + 2]
            }
            for (int c = te; c < N; c += kWarps * 32) {
                float val[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float* px = prow[u] + c;
                    float ss;
                    if (win5) {
                        const float a0 = px[0], a1 = px[1], a2 = px[2], a3 = px[3], a4 = px[4];
                        ss = fmaf(a0, a0, fmaf(a1, a1, fmaf(a2, a2, fmaf(a3, a3, a4 * a4))));
                    } else {
                        ss = 0.f;
                        const int c_lo = max(0, c - half), c_hi = min(N - 1, c + half);
                        for (int qq = c_lo; qq <= c_hi; ++qq) ss = fmaf(prow[u][qq + 2], prow[u][qq + 2], ss);
                    }
                    // x / (k + a*ss)^beta; the base is >= k > 0, so exp2/log2 is safe
                    val[u] = px[2] * exp2f(nbeta * __log2f(g.lrn.k + al * ss));
                    if (p.round_out) val[u] = to_tf32(val[u]);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (optr[u]) optr[u][c] = val[u];
            }
        }
    }
    bar_group<kBar, kWarps>();                                // cbuf / pbuf free again, all stores issued
    if (g.n_signal > 0 && te == 0) tile_signal(g, n, pr0, min(pr0 + p.PR, g.p1));   // producer side of the halo flags
}

// pooled-pixel table (cbuf offset of the window origin, (pi,pj)) and the zero borders of the LRN rows
template <int kThreadsInGroup>
__device__ __forceinline__ void epilogue_tables(const TcParams& p, int2* ptab, float* pbuf, int te) {
    const int PP = p.PR * p.PC, N = p.g.K, NP = N + 4;
    if (te < PP) {
        const int pi = te / p.PC, pj = te - pi * p.PC;
        ptab[te] = make_int2(((2 * pi) * p.TC + 2 * pj) * kCbufPitch, (pi << 16) | pj);
    }
    if (p.g.lrn.N_lrn > 0)
        for (int i = te; i < PP * 4; i += kThreadsInGroup)
            pbuf[(i >> 2) * NP + ((i & 3) < 2 ? (i & 3) : N + (i & 3))] = 0.f;
}

// ------------------------------------------------------------------------ kernel ----
// kPatch = false ("vec" mode, C % 32 == 0: conv2..5): a k-block is 32 channels of one filter tap;
//   every A row is 128 contiguous bytes of the activation -> cp.async 16 B straight into the
//   swizzled stage, completion tracked by the stage's mbarrier (cp.async.mbarrier.arrive.noinc).
//   K order (channel-chunk, fy, fx, c%32): consecutive k-blocks re-read neighbouring pixels, so
//   the small L1 that is left catches most of the im2col re-reads (local and NVLink-mapped rows).
// kPatch = true  (conv1, C = 3): the input patch of the tile is staged ONCE in shared memory (one
//   pull over NVLink in p2p-scatter mode instead of (F/S)^2 re-reads), then expanded to A with
//   LDS.128 -> cvt.rna.tf32 -> STS.128.  K order (fy, [fx,c] padded to a multiple of 4): for a fixed
//   filter row the F*C inputs of a pixel are contiguous, so every 16 B chunk is an aligned LDS.128.
template <bool kPatch>
__global__ void __launch_bounds__(kThreads, 1)
tc_block_kernel(const __grid_constant__ CUtensorMap tmapB, const __grid_constant__ TcParams p,
                const float* __restrict__ bias, float* __restrict__ out) {
    extern __shared__ uint8_t smem_raw[];
    const b200_block_geom& g = p.g;
    const int N = g.K;                                        // GEMM N = output channels

    // carve shared memory (1024 B aligned for the swizzle atoms)
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t b_off = p.stages * kAStageBytes;
    const uint32_t full_bar = base + p.ctl_off;               // [stages]
    const uint32_t empty_bar = full_bar + 8 * p.stages;       // [stages]
    const uint32_t tmem_full_bar = empty_bar + 8 * p.stages;
    const uint32_t tile_free_bar = tmem_full_bar + 8;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gbase + p.ctl_off + 16 * p.stages + 16);
    const float** rowtab = reinterpret_cast<const float**>(gbase + p.ctl_off + 16 * p.stages + 32);
    float* s_bias = reinterpret_cast<float*>(gbase + p.ctl_off + 16 * p.stages + 32 + kRowTab * 8 +
                                             kMaxPooled * 8);
    int2* ptab = reinterpret_cast<int2*>(gbase + p.ctl_off + 16 * p.stages + 32 + kRowTab * 8);
    float* patch = reinterpret_cast<float*>(gbase + p.patch_off);
    float* cbuf0 = reinterpret_cast<float*>(gbase);           // epilogue buffers alias the stages
    float* pbuf = reinterpret_cast<float*>(gbase + 2 * kCbufBytes);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) {
            // vec mode: every producer thread's cp.async group arrives (noinc); patch mode: one per warp
            mbar_init(full_bar + 8 * s, (kPatch ? kProducerWarps : kProducerThreads) + 1);
            mbar_init(empty_bar + 8 * s, 1);
        }
        mbar_init(tmem_full_bar, 1);
        mbar_init(tile_free_bar, kProducerWarps);
        fence_mbar_init();
        block_gate_wait(g);
    }
    for (int i = threadIdx.x; i < N; i += kThreads) s_bias[i] = bias[i];
    if (warp == kMmaWarp) tmem_alloc(smem_u32(tmem_slot), p.tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int tiles_per_img = p.tiles_r * p.tiles_c;
    int stage = 0;
    uint32_t phase = 0;

    if (warp < kProducerWarps) {
        // =============================== producers, then epilogue =========================
        const int tid = threadIdx.x;
        const int j = tid & 7;                                // 16 B chunk inside the 128 B row
        const int m_base = tid >> 3;                          // rows m_base + 32*i
        const uint32_t row_sw = (uint32_t)(m_base & 7);       // (m & 7) is the same for all i
        // byte offset of (row m_base, chunk j) inside a stage; row m_base+32*i adds i*4096
        const uint32_t dst0 = (uint32_t)(m_base >> 3) * 1024u + row_sw * 128u + (((uint32_t)j ^ row_sw) << 4);
        const int IRT = (p.TR - 1) * g.S + g.F;               // input rows a tile touches
        int pij[8];                                           // (pi << 16 | pj) of row i, -1 = unused slot
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int m = m_base + 32 * i;
            const int pi = m / p.TC;
            pij[i] = m < p.TR * p.TC ? (pi << 16) | (m - pi * p.TC) : -1;
        }
        int it = 0;
        for (long long w = blockIdx.x; w < p.total_tiles; w += gridDim.x, ++it) {
            const int n = (int)(w / tiles_per_img);
            const int rem = (int)(w - (long long)n * tiles_per_img);
            const int tr_i = rem / p.tiles_c, tc_i = rem - tr_i * p.tiles_c;
            const int pr0 = g.p0 + tr_i * p.PR, pc0 = tc_i * p.PC;
            const int cr0 = 2 * pr0, cc0 = 2 * pc0;
            const int ir0 = cr0 * g.S - g.P;                   // first input row of the tile

            if (p.any_flag) {                                  // consumer side of the halo flags
                if (tid == 0) {
                    for (int s = 0; s < g.nseg; ++s)
                        if (g.seg[s].flag && ir0 < g.seg[s].hi && ir0 + IRT > g.seg[s].lo)
                            wait_flag_ge(g.seg[s].flag + n, g.seg[s].target * (g.epoch ? *g.epoch : 1), g.status, g.timeout_ns);
                }
                bar_producers();
            }

            // per-thread description of its 8 A rows (pixels m_base + 32*i)
            int rsel[8];        // vec: row-table index of the pixel's first input row; patch: float offset
            int cbyte[8];       // vec: byte offset of (first tap column, chunk j) inside an input row
            uint32_t xmask[8];  // vec: bit fx set <=> input column x0+fx lies inside the image
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int pi = pij[i] >> 16, pj = pij[i] & 0xFFFF;
                const bool ok = pij[i] >= 0 && (cr0 + pi) < g.Hc && (cc0 + pj) < g.Wc;
                if (kPatch) {
                    rsel[i] = ok ? pi * g.S * p.patch_pitch + pj * g.S * g.C : 0;
                    cbyte[i] = 0; xmask[i] = 0;
                } else {
                    const int x0 = (cc0 + pj) * g.S - g.P;
                    rsel[i] = ok ? pi * g.S : IRT;             // IRT.. = null rows -> zero fill
                    cbyte[i] = (x0 * g.C + 4 * j) * 4;
                    const int lo = max(0, -x0), hi = min(g.F, g.W - x0);   // valid taps [lo,hi)
                    xmask[i] = hi > lo ? (((1u << hi) - 1u) & ~((1u << lo) - 1u)) : 0u;
                }
            }

            // ---- row-pointer table: base pointer of every input row this tile touches ----
            {
                const int nrows = kPatch ? p.patch_rows : IRT + g.F;
                const int nreal = kPatch ? p.patch_rows : IRT;
                if (tid < nrows) {
                    const int h = ir0 + tid;
                    rowtab[tid] = (tid < nreal && h >= 0 && h < g.H) ? seg_row_ptr(g, n, h) : nullptr;
                }
                bar_producers();
            }

            if (kPatch) {
                // ---- stage the input patch once (rows ir0.., float columns cf0..) ----
                const int cf0 = (cc0 * g.S - g.P) * g.C, WC = g.W * g.C;
                const uint32_t patch_s = smem_u32(patch);
                for (int f = tid; f < p.patch_pitch; f += kProducerThreads) {
                    const int cf = cf0 + f;
                    const bool cv = cf >= 0 && cf < WC;
                    uint32_t dst = patch_s + (uint32_t)f * 4u;
                    for (int r = 0; r < p.patch_rows; ++r, dst += (uint32_t)p.patch_pitch * 4u) {
                        const float* rp = rowtab[r];
                        const bool v = cv && rp != nullptr;
                        cp_async4(dst, rp + cf, v ? 4u : 0u);   // size 0: nothing is read, zero fill
                    }
                }
                cp_async_wait_all();
                bar_producers();
            }

            // k-block walk state
            int fy = 0, fx = 0, cch = 0;                        // vec: tap + channel chunk (uniform)
            int pfy = (4 * j) / p.FCpad, pt = (4 * j) - pfy * p.FCpad;   // patch: this thread's chunk
            const char* rp8[8];                                 // vec: row pointers for the current fy
            uint32_t vm[8];                                     // vec: xmask, or 0 when the row is padding
            for (int kb = 0; kb < ((p.dbg & 8) ? 0 : p.nkb); ++kb) {
                if (!kPatch && fx == 0) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        rp8[i] = reinterpret_cast<const char*>(rowtab[rsel[i] + fy]);
                        vm[i] = rp8[i] ? xmask[i] : 0u;
                    }
                }
                mbar_wait(empty_bar + 8 * stage, phase ^ 1);
                const uint32_t a_dst = base + stage * kAStageBytes + dst0;
                const uint32_t fb = full_bar + 8 * stage;
                if (p.dbg & 1) {
                    if (kPatch) { __syncwarp(); if (lane == 0) mbar_arrive(fb); } else mbar_arrive(fb);
                } else if (kPatch) {
                    const float* src = patch + pfy * p.patch_pitch + pt;
                    float4 v[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) v[i] = *reinterpret_cast<const float4*>(src + rsel[i]);
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a_dst + i * 4096u),
                                     "f"(to_tf32(v[i].x)), "f"(to_tf32(v[i].y)), "f"(to_tf32(v[i].z)),
                                     "f"(to_tf32(v[i].w))
                                     : "memory");
                    fence_proxy_async_smem();                  // generic stores -> async proxy (UMMA)
                    __syncwarp();
                    if (lane == 0) mbar_arrive(fb);            // one arrival per warp (256 would serialise)
                    pt += kBlockK;
                    while (pt >= p.FCpad) { pt -= p.FCpad; ++pfy; }
                } else {
                    const int off = (fx * g.C + cch * kBlockK) * 4;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const uint32_t nb = ((vm[i] >> fx) & 1u) ? 16u : 0u;   // 0 => zero fill, no read
                        cp_async16(a_dst + i * 4096u, rp8[i] + (cbyte[i] + off), nb);
                    }
                    cp_async_mbar_arrive_noinc(fb);            // arrives when this thread's copies land
                }
                if (!kPatch) {
                    if (++fx == g.F) { fx = 0; if (++fy == g.F) { fy = 0; ++cch; } }
                }
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
            // ------------------------------ epilogue ------------------------------------
            mbar_wait(tmem_full_bar, (uint32_t)(it & 1));
            tc_fence_after();
            epilogue_tables<kProducerThreads>(p, ptab, pbuf, tid);   // pbuf aliases the stages: borders per tile
            epilogue_pool_lrn<1, kProducerWarps>(p, tmem_base, cbuf0, kCbufBytes / 4, pbuf, ptab, s_bias, out, n, pr0,
                                                 pc0, warp, lane, (warp >> 2) < p.nacc, [] {});
            if (lane == 0) mbar_arrive(tile_free_bar);
        }
    } else if (warp == kMmaWarp) {
        // ======================================= MMA issuer ==============================
        for (long long w = blockIdx.x; w < p.total_tiles; w += gridDim.x) {
            for (int kb = 0; kb < p.nkb; ++kb) {
                if (!(p.dbg & 8)) mbar_wait(full_bar + 8 * stage, phase);
                fence_proxy_async_smem();                       // cp.async (generic proxy) wrote A; the MMA reads it
                tc_fence_after();                               // through the async proxy
                if (lane == 0) {
                    const uint32_t a_stage = base + stage * kAStageBytes;
                    const uint32_t b_stage = base + b_off + stage * p.b_stage_bytes;
                    const uint64_t db = make_smem_desc(b_stage);
#pragma unroll
                    for (int ks = 0; ks < kBlockK / 8; ++ks) {
                        for (int h = 0; h < p.nacc; ++h) {
                            const uint64_t da = make_smem_desc(a_stage + h * (128 * 128));
                            umma_tf32(tmem_base + (uint32_t)(h * N), da + 2 * ks, db + 2 * ks, p.idesc,
                                      (kb | ks) != 0 ? 1u : 0u);
                        }
                    }
                    if (!(p.dbg & 8)) umma_commit(empty_bar + 8 * stage);   // frees the stage when the MMAs retire
                    if (kb == p.nkb - 1) umma_commit(tmem_full_bar);
                }
                __syncwarp();
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        }
    } else {
        // ======================================= TMA issuer (B) ==========================
        int it = 0;
        for (long long w = blockIdx.x; w < p.total_tiles; w += gridDim.x, ++it) {
            // the previous tile's epilogue buffers alias the stages: wait until they are released
            if (it > 0) mbar_wait(tile_free_bar, (uint32_t)((it - 1) & 1));
            for (int kb = 0; kb < ((p.dbg & 8) ? 0 : p.nkb); ++kb) {
                mbar_wait(empty_bar + 8 * stage, phase ^ 1);
                if (lane == 0) {
                    const uint32_t fb = full_bar + 8 * stage;
                    if (p.dbg & 2) {
                        mbar_arrive(fb);
                    } else {
                        mbar_arrive_expect_tx(fb, (uint32_t)p.b_stage_bytes);
                        tma_load_2d(base + b_off + stage * p.b_stage_bytes, &tmapB, kb * kBlockK, 0, fb);
                    }
                }
                __syncwarp();
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after();
        tmem_dealloc(tmem_base, p.tmem_cols);
    }
    if (threadIdx.x == 0) block_done_signal(g);
}

// ------------------------------------------------------- small-K_out kernel (conv1) ----
// tc_block_kernel spends ~160 cycles per tcgen05.mma whatever N is (profiles/tc_experiments_v5.log:
// with N=96 the tensor pipe is instruction-bound, not FLOP-bound), and its producer warps double as
// epilogue warps, so staging, MMA and epilogue of a tile run one after the other.  For layers with
// few output channels (K_out <= 128: conv1) this kernel turns the GEMM around and splits the roles:
//   * A operand = weights [K_out -> 128 rows][K] (TMA), B operand = the im2col pixels, N = 128/256
//     pixel slots per instruction: half as many tcgen05.mma per tile;
//   * D is channels x pixels: 128 TMEM lanes x 256 columns = half of TMEM, so accumulators are
//     double buffered and tile t's epilogue overlaps tile t+1's staging and MMAs;
//   * warps 0-7 only stage the input patch and build B; warp 8 issues MMAs; warp 9 issues TMA;
//     warps 10-13 are the epilogue: warp q owns TMEM lane quarter q = channels 32q..32q+31, drains
//     three conv rows at a time through a private smem strip, max-pools and stores — no CTA-wide
//     barrier anywhere in the epilogue.
constexpr int kSwThreads = 576;                     // 8 producer + MMA + TMA + 8 epilogue warps
constexpr int kSwEpiWarps = 8;                      // two per TMEM lane quarter: even / odd pooled rows
constexpr int kSwCtlHead = 208;                      // barriers + TMEM slot in front of the row table
constexpr int kSwCtlBytes = kSwCtlHead + kRowTab * 8 + 128 * 4;
constexpr int kSwMaxStages = 6;

struct SwParams {
    TcParams t;
    int nload;                                       // 32-column TMEM loads per pooled row (3*TC slots)
    uint32_t strip_off, strip_bytes;                 // per-epilogue-warp smem strip
    uint32_t sw_ctl_off;
    uint32_t w_stage_bytes;                          // one weight slot: K_out rows x 128 B, whole 1 KB swizzle atoms
    uint32_t sw_idesc;
    int direct;                                      // 1: B rows are cp.async'd straight from global memory
    int nstages;                                     // 2 (patch) or 3 (direct)
    int cpr;                                         // direct: 16 B chunks per filter row run (F*C/4)
    int nkb_main;                                    // pixel-TMA mode: k-blocks of whole filter rows (= F), rest = tails
    uint32_t pix_bytes;                              // pixel-TMA mode: bytes of one pixel box (32 x TC x TR floats)
    int regpool;                                     // epilogue pools in registers (TC <= 16): no smem strips
};

// Pixel-TMA mode (the default conv1 loader; B200_TC_CONV1=cpasync selects the other one): one tensor map per input
// row phase r = fy % S over an overlapping-stride view of the resident input,
//   element (i, x, y, n) = in[n][S*y + r][x*S*C + i],  dims {32, Wc, rows/S, N},
// so the first 32 floats of every pixel's filter-row run, for a whole TR x TC tile, are ONE box {32, TC, TR, 1}
// that lands in the 128B-swizzled stage in pixel-slot order.
struct alignas(64) PixMaps {
    CUtensorMap m[4];
};

template <bool kPixTma>
__global__ void __launch_bounds__(kSwThreads, 1)
tc_block_swapped_kernel(const __grid_constant__ CUtensorMap tmapW, const __grid_constant__ SwParams sp,
                        const float* __restrict__ bias, float* __restrict__ out,
                        const __grid_constant__ PixMaps pm) {
    extern __shared__ uint8_t smem_raw[];
    const TcParams& p = sp.t;
    const b200_block_geom& g = p.g;
    const int KO = g.K;                                       // output channels (<= 128)
    const int slots = 128 * p.nacc;                           // GEMM N = pixel slots

    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
    const int NS = sp.nstages;
    const uint32_t wsb = sp.w_stage_bytes;                    // weight slots first (K_out rows x 128 B each), then the
    const uint32_t pixs = base + NS * wsb;                     // pixel stages: the MMA's rows K_out..127 read on into smem we own
    const uint32_t full_bar = base + sp.sw_ctl_off;           // [8]
    const uint32_t empty_bar = full_bar + 64;                 // [8]
    const uint32_t tfull_bar = full_bar + 128;                // [2] accumulator ready
    const uint32_t tempty_bar = full_bar + 144;               // [2] accumulator drained
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gbase + sp.sw_ctl_off + 160);
    const float** rowtab = reinterpret_cast<const float**>(gbase + sp.sw_ctl_off + kSwCtlHead);
    float* s_bias = reinterpret_cast<float*>(gbase + sp.sw_ctl_off + kSwCtlHead + kRowTab * 8);
    // Pixel-TMA mode, tail k-block t (0 or 1) of every tile: the producers cp.async its pixel half into the ring stage
    // the block will use.  tail_free[t]: that stage's previous occupant (k-block nkb_main + t - NS of the same tile) has
    // been consumed; tail_bar[t]: the copies have landed.  Both complete once per tile, so their parity is the tile's;
    // the ring's own empty barriers cannot be used here because the producers skip the main blocks and a parity wait
    // is only safe for a waiter that follows every phase.
    const uint32_t tail_bar = full_bar + 176;                 // [2]
    const uint32_t tail_free = full_bar + 192;                // [2]
    float* patch = reinterpret_cast<float*>(gbase + p.patch_off);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; ++s) {
            // pixel-TMA mode: the TMA warp's expect_tx is the only arrival (tail k-blocks: see tail_bar)
            mbar_init(full_bar + 8 * s, kPixTma ? 1 : (sp.direct ? kProducerThreads : kProducerWarps) + 1);
            mbar_init(empty_bar + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(tfull_bar + 8 * b, 1);
            mbar_init(tempty_bar + 8 * b, kSwEpiWarps);
            if (kPixTma) {
                mbar_init(tail_bar + 8 * b, kProducerThreads);     // one asynchronous arrival per producer thread
                mbar_init(tail_free + 8 * b, 1);                   // tcgen05.commit by the MMA warp
            }
        }
        fence_mbar_init();
        block_gate_wait(g);
    }
    for (int i = threadIdx.x; i < 128; i += kSwThreads) s_bias[i] = i < KO ? bias[i] : 0.f;
    if (warp == kMmaWarp) tmem_alloc(smem_u32(tmem_slot), 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    grid_dep_launch();          // the next block's CTAs may take SMs as ours exit (they wait for this grid's end)
    const uint32_t tmem_base = *tmem_slot;
    const int tiles_per_img = p.tiles_r * p.tiles_c;

    if (warp < kProducerWarps) {
        // ============================ producers: patch staging + im2col (B operand) ==========
        const int tid = threadIdx.x;
        const int j = tid & 7, m_base = tid >> 3;
        const uint32_t row_sw = (uint32_t)(m_base & 7);
        const uint32_t dst0 = (uint32_t)(m_base >> 3) * 1024u + row_sw * 128u + (((uint32_t)j ^ row_sw) << 4);
        const int IRT = (p.TR - 1) * g.S + g.F;
        int pij[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int m = m_base + 32 * i;
            const int pi = m / p.TC;
            pij[i] = m < p.TR * p.TC ? (pi << 16) | (m - pi * p.TC) : -1;
        }
        int stage = 0;
        uint32_t phase = 0;
        long long tile_it = 0;
        for (long long w = blockIdx.x; w < p.total_tiles; w += gridDim.x, ++tile_it) {
            const int n = (int)(w / tiles_per_img);
            const int rem = (int)(w - (long long)n * tiles_per_img);
            const int tr_i = rem / p.tiles_c, tc_i = rem - tr_i * p.tiles_c;
            const int cr0 = 2 * (g.p0 + tr_i * p.PR), cc0 = 2 * (tc_i * p.PC);
            const int ir0 = cr0 * g.S - g.P;
            if (p.any_flag) {
                if (tid == 0)
                    for (int s = 0; s < g.nseg; ++s)
                        if (g.seg[s].flag && ir0 < g.seg[s].hi && ir0 + IRT > g.seg[s].lo)
                            wait_flag_ge(g.seg[s].flag + n, g.seg[s].target * (g.epoch ? *g.epoch : 1),
                                         g.status, g.timeout_ns);
            }
            int rsel[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int pi = pij[i] >> 16, pj = pij[i] & 0xFFFF;
                const bool ok = pij[i] >= 0 && (cr0 + pi) < g.Hc && (cc0 + pj) < g.Wc;
                rsel[i] = ok ? pi * g.S * p.patch_pitch + pj * g.S * g.C : 0;
            }
            bar_producers();                                   // previous tile's loads/im2col are issued
            if (sp.direct) {
                // ---- direct mode: rows of B are 16 B chunks of the activation itself ----
                if (tid < IRT) {
                    const int h = ir0 + tid;
                    rowtab[tid] = (h >= 0 && h < g.H) ? seg_row_ptr(g, n, h) : nullptr;
                }
                bar_producers();
                int hsel[8], xb[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int pi = pij[i] >> 16, pj = pij[i] & 0xFFFF;
                    const bool ok = pij[i] >= 0 && (cr0 + pi) < g.Hc && (cc0 + pj) < g.Wc;
                    hsel[i] = ok ? pi * g.S : -1;
                    xb[i] = ((cc0 + pj) * g.S - g.P) * g.C * 4;
                }
                if (sp.direct == 2 && kPixTma) {
                    // Row K order, main k-blocks by TMA (warp 9): only the tail blocks are ours.  Their ring position
                    // follows from the k-block count; the copies are fire-and-forget (each thread's cp.asyncs arrive
                    // on the tail barrier when they land), so a producer never waits for data.
                    for (int kb = sp.nkb_main; kb < p.nkb; ++kb) {
                        const int tt = kb - sp.nkb_main;
                        const int st = (int)((tile_it * p.nkb + kb) % NS);
                        mbar_wait(tail_free + 8 * tt, (uint32_t)(tile_it & 1));
                        const uint32_t a_dst = pixs + st * kAStageBytes + dst0;
                        const int tfy = 8 * tt + j;
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            if (i >= 4 * p.nacc) break;
                            const char* rp = (tfy < g.F && hsel[i] >= 0)
                                                 ? reinterpret_cast<const char*>(rowtab[hsel[i] + tfy]) : nullptr;
                            if (!(p.dbg & 4096)) cp_async16(a_dst + i * 4096u, rp + (xb[i] + 128), rp ? 16u : 0u);
                        }
                        cp_async_mbar_arrive_noinc(tail_bar + 8 * tt);
                    }
                    continue;
                }
                if (sp.direct == 2) {
                    // Row K order (shared by both loaders, so they are bit-identical): k-block kb < nkb_main holds
                    // floats 0..31 of filter row kb's run of every pixel; tail block t holds, as chunk j, the 16 B at
                    // run offset 128 B of filter row 8t + j.
                    for (int kb = 0; kb < p.nkb; ++kb) {
                        mbar_wait(empty_bar + 8 * stage, phase ^ 1);
                        const bool tail = kb >= sp.nkb_main;
                        const uint32_t a_dst = pixs + stage * kAStageBytes + dst0;
                        const int tfy = tail ? 8 * (kb - sp.nkb_main) + j : kb;
                        const int boff = tail ? 128 : 16 * j;
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            if (i >= 4 * p.nacc) break;
                            const char* rp = (tfy < g.F && hsel[i] >= 0)
                                                 ? reinterpret_cast<const char*>(rowtab[hsel[i] + tfy]) : nullptr;
                            cp_async16(a_dst + i * 4096u, rp + (xb[i] + boff), rp ? 16u : 0u);
                        }
                        cp_async_mbar_arrive_noinc(full_bar + 8 * stage);
                        if (++stage == NS) { stage = 0; phase ^= 1; }
                    }
                    continue;
                }
                int fy = 0, ch = j;                             // this thread's chunk: run fy, chunk ch in run
                while (ch >= sp.cpr) { ch -= sp.cpr; ++fy; }
                for (int kb = 0; kb < p.nkb; ++kb) {
                    mbar_wait(empty_bar + 8 * stage, phase ^ 1);
                    const uint32_t a_dst = pixs + stage * kAStageBytes + dst0;
                    const bool kv = fy < g.F;                   // beyond F*F*C: zero columns of the K padding
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        if (i >= 4 * p.nacc) break;
                        const char* rp = (kv && hsel[i] >= 0) ? reinterpret_cast<const char*>(rowtab[hsel[i] + fy])
                                                              : nullptr;
                        cp_async16(a_dst + i * 4096u, rp + (xb[i] + ch * 16), rp ? 16u : 0u);
                    }
                    cp_async_mbar_arrive_noinc(full_bar + 8 * stage);
                    ch += 8;
                    while (ch >= sp.cpr) { ch -= sp.cpr; ++fy; }
                    if (++stage == NS) { stage = 0; phase ^= 1; }
                }
                continue;
            }
            if (tid < p.patch_rows) {
                const int h = ir0 + tid;
                rowtab[tid] = (h >= 0 && h < g.H) ? seg_row_ptr(g, n, h) : nullptr;
            }
            bar_producers();
            {
                const int cf0 = (cc0 * g.S - g.P) * g.C, WC = g.W * g.C;
                const uint32_t patch_s = smem_u32(patch);
                for (int f = tid; f < p.patch_pitch; f += kProducerThreads) {
                    const int cf = cf0 + f;
                    const bool cv = cf >= 0 && cf < WC;
                    uint32_t dst = patch_s + (uint32_t)f * 4u;
                    for (int r = 0; r < p.patch_rows; ++r, dst += (uint32_t)p.patch_pitch * 4u) {
                        const float* rp = rowtab[r];
                        cp_async4(dst, rp + cf, (cv && rp != nullptr) ? 4u : 0u);
                    }
                }
                cp_async_wait_all();
                bar_producers();
            }
            int pfy = (4 * j) / p.FCpad, pt = (4 * j) - pfy * p.FCpad;
            for (int kb = 0; kb < p.nkb; ++kb) {
                mbar_wait(empty_bar + 8 * stage, phase ^ 1);
                const uint32_t a_dst = pixs + stage * kAStageBytes + dst0;
                const float* src = patch + pfy * p.patch_pitch + pt;
                float4 v[8];                                    // all loads first: the stores are asm
#pragma unroll                                                  // volatile and would fence each load
                for (int i = 0; i < 8; ++i) v[i] = *reinterpret_cast<const float4*>(src + rsel[i]);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    if (i >= 4 * p.nacc) break;                 // nacc == 1: only 128 pixel slots
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a_dst + i * 4096u),
                                 "f"(to_tf32(v[i].x)), "f"(to_tf32(v[i].y)), "f"(to_tf32(v[i].z)),
                                 "f"(to_tf32(v[i].w))
                                 : "memory");
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(full_bar + 8 * stage);
                pt += kBlockK;
                while (pt >= p.FCpad) { pt -= p.FCpad; ++pfy; }
                if (++stage == NS) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == kMmaWarp) {
        // ======================================= MMA issuer ==============================
        int stage = 0, buf = 0;
        uint32_t phase = 0, bphase = 0, tile_par = 0;
        for (long long w = blockIdx.x; w < p.total_tiles; w += gridDim.x, tile_par ^= 1u) {
            mbar_wait(tempty_bar + 8 * buf, bphase ^ 1);        // epilogue has drained this accumulator
            tc_fence_after();
            for (int kb = 0; kb < p.nkb; ++kb) {
                mbar_wait(full_bar + 8 * stage, phase);
                if (kPixTma && kb >= sp.nkb_main)               // pixel half of a tail block: the producers' cp.asyncs
                    mbar_wait(tail_bar + 8 * (kb - sp.nkb_main), tile_par);
                fence_proxy_async_smem();                       // producers' cp.async data -> async-proxy reads
                tc_fence_after();
                if (lane == 0) {
                    const uint64_t dpix = make_smem_desc(pixs + stage * kAStageBytes);
                    const uint64_t dwgt = make_smem_desc(base + stage * wsb);
                    if (!(p.dbg & 512)) {
#pragma unroll
                        for (int ks = 0; ks < kBlockK / 8; ++ks)
                            umma_tf32(tmem_base + (uint32_t)(buf * 256), dwgt + 2 * ks, dpix + 2 * ks, sp.sw_idesc,
                                      (kb | ks) != 0 ? 1u : 0u);
                    }
                    umma_commit(empty_bar + 8 * stage);
                    if (kPixTma && kb + NS >= sp.nkb_main && kb + NS < p.nkb)      // the next user of this stage is a tail block
                        umma_commit(tail_free + 8 * (kb + NS - sp.nkb_main));
                    if (kb == p.nkb - 1) umma_commit(tfull_bar + 8 * buf);
                }
                __syncwarp();
                if (++stage == NS) { stage = 0; phase ^= 1; }
            }
            if (++buf == 2) { buf = 0; bphase ^= 1; }
        }
    } else if (warp == kTmaWarp) {
        // ======================================= TMA issuer (weights) ====================
        int stage = 0;
        uint32_t phase = 0;
        for (long long w = blockIdx.x; w < p.total_tiles; w += gridDim.x) {
            int pn = 0, pcr0 = 0, pcc0 = 0;                     // pixel-TMA mode: image, conv-row/col origin of the tile
            if (kPixTma) {
                pn = (int)(w / tiles_per_img);
                const int rem = (int)(w - (long long)pn * tiles_per_img);
                const int tr_i = rem / p.tiles_c, tc_i = rem - tr_i * p.tiles_c;
                pcr0 = 2 * (g.p0 + tr_i * p.PR);
                pcc0 = 2 * (tc_i * p.PC);
            }
            for (int kb = 0; kb < p.nkb; ++kb) {
                mbar_wait(empty_bar + 8 * stage, phase ^ 1);
                if (lane == 0) {
                    const uint32_t fb = full_bar + 8 * stage;
                    // dbg 64 / 128: timing experiments without the pixel / weight loads (results are garbage)
                    const bool pix = kPixTma && kb < sp.nkb_main && !(p.dbg & 64);
                    const bool wgt = !(p.dbg & 128);
                    mbar_arrive_expect_tx(fb, (wgt ? (uint32_t)(KO * 128) : 0u) + (pix ? sp.pix_bytes : 0u));
                    if (pix)    // filter row fy = kb: input row S*(cr0+pi) + fy = S*(cr0 + pi + fy/S) + fy%S
                        tma_load_4d(pixs + stage * kAStageBytes, &pm.m[kb % g.S], 0, pcc0,
                                    pcr0 + kb / g.S - g.seg[0].base / g.S, pn, fb);
                    if (wgt) tma_load_2d(base + stage * wsb, &tmapW, kb * kBlockK, 0, fb);
                }
                __syncwarp();
                if (++stage == NS) { stage = 0; phase ^= 1; }
            }
        }
    } else {
        // ======================================= epilogue warps ==========================
        const int q = warp & 3;                                // a warp can only touch TMEM lane quarter warp%4;
                                                               // quarter q == channels 32q..32q+31
        const int half = (warp - 10) >> 2;                     // two warps per quarter: even / odd pooled rows
        const bool live = q * 32 < KO;
        const int ch = q * 32 + lane;
        float* strip = reinterpret_cast<float*>(gbase + sp.strip_off + (half * 4 + q) * sp.strip_bytes);
        const float bch = s_bias[ch];
        int buf = 0;
        uint32_t bphase = 0;
        for (long long w = blockIdx.x; w < p.total_tiles; w += gridDim.x) {
            const int n = (int)(w / tiles_per_img);
            const int rem = (int)(w - (long long)n * tiles_per_img);
            const int tr_i = rem / p.tiles_c, tc_i = rem - tr_i * p.tiles_c;
            const int pr0 = g.p0 + tr_i * p.PR, pc0 = tc_i * p.PC;
            mbar_wait(tfull_bar + 8 * buf, bphase);
            tc_fence_after();
            if (live && sp.regpool && !(p.dbg & 256)) {
                // Lane = channel and TMEM columns = pixel slots, so a 3x3 window is nine registers of ONE lane:
                // three 16-column loads (conv rows 2pr, 2pr+1, 2pr+2 of the tile; TC <= 16) and the row is pooled
                // without touching shared memory.  The stores are 128 B per pixel (32 channels of the warp).
                for (int pr = half; pr < p.PR; pr += 2) {
                    const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * 256 + 2 * pr * p.TC);
                    uint32_t ra[16], rb[16], rc[16];
                    tmem_ld16(t0, ra);
                    tmem_ld16(t0 + (uint32_t)p.TC, rb);
                    tmem_ld16(t0 + (uint32_t)(2 * p.TC), rc);
                    const int prg = pr0 + pr;
                    const bool row_ok = prg < g.p1 && prg < g.Hp && ch < KO;
                    float* orow = out + (((size_t)n * g.out_slab + (prg - g.out_base)) * g.Wp) * KO + ch;
                    float* mrow[B200_MAX_MIRROR] = {nullptr, nullptr, nullptr, nullptr};
                    bool pushed = false;
                    if (g.n_mirror > 0) {
#pragma unroll
                        for (int m = 0; m < B200_MAX_MIRROR; ++m) {
                            const bool in = m < g.n_mirror && prg >= g.mirror[m].lo && prg < g.mirror[m].hi;
                            if (in)
                                mrow[m] = g.mirror[m].ptr +
                                          (((size_t)n * g.mirror[m].slab + (prg - g.mirror[m].base)) * g.Wp) * KO + ch;
                            pushed |= in;
                        }
                    }
#pragma unroll
                    for (int pc = 0; pc < 7; ++pc) {
                        if (pc < p.PC) {
                            float mx = fmaxf(fmaxf(__uint_as_float(ra[2 * pc]), __uint_as_float(ra[2 * pc + 1])),
                                             __uint_as_float(ra[2 * pc + 2]));
                            mx = fmaxf(mx, fmaxf(fmaxf(__uint_as_float(rb[2 * pc]), __uint_as_float(rb[2 * pc + 1])),
                                                 __uint_as_float(rb[2 * pc + 2])));
                            mx = fmaxf(mx, fmaxf(fmaxf(__uint_as_float(rc[2 * pc]), __uint_as_float(rc[2 * pc + 1])),
                                                 __uint_as_float(rc[2 * pc + 2])));
                            float val = fmaxf(mx + bch, 0.f);
                            if (p.round_out) val = to_tf32(val);
                            const int pcg = pc0 + pc;
                            if (row_ok && pcg < g.Wp) {
                                orow[(size_t)pcg * KO] = val;
                                if (pushed) {
#pragma unroll
                                    for (int m = 0; m < B200_MAX_MIRROR; ++m)
                                        if (mrow[m]) mrow[m][(size_t)pcg * KO] = val;
                                }
                            }
                        }
                    }
                }
            } else if (live && !(p.dbg & 256)) {               // dbg 256: timing experiment without the epilogue's work
                for (int pr = half; pr < p.PR; pr += 2) {
                    // conv rows 2pr..2pr+2 = slots [col0, col0 + 3*TC); load whole 32-column groups
                    const int col0 = 2 * pr * p.TC;
                    int start = col0;
                    if (start + 32 * sp.nload > slots) start = slots - 32 * sp.nload;
                    const int shift = col0 - start;
                    for (int l = 0; l < sp.nload; ++l) {
                        uint32_t r[32];
                        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * 256 + start + 32 * l), r);
#pragma unroll
                        for (int i = 0; i < 32; ++i) strip[(32 * l + i) * kCbufPitch + lane] = __uint_as_float(r[i]);
                    }
                    __syncwarp();
                    const int prg = pr0 + pr;
                    const bool row_ok = prg < g.p1 && prg < g.Hp && ch < KO;
                    float* orow = out + (((size_t)n * g.out_slab + (prg - g.out_base)) * g.Wp) * KO + ch;
                    // halo push: this row's copy in the neighbours' slabs (nullptr: not read by that neighbour)
                    float* mrow[B200_MAX_MIRROR] = {nullptr, nullptr, nullptr, nullptr};
                    bool pushed = false;
                    if (g.n_mirror > 0) {                       // kernel-uniform: single-GPU launches skip all of it
#pragma unroll
                        for (int m = 0; m < B200_MAX_MIRROR; ++m) {
                            const bool in = m < g.n_mirror && prg >= g.mirror[m].lo && prg < g.mirror[m].hi;
                            if (in)
                                mrow[m] = g.mirror[m].ptr +
                                          (((size_t)n * g.mirror[m].slab + (prg - g.mirror[m].base)) * g.Wp) * KO + ch;
                            pushed |= in;
                        }
                    }
                    for (int pc = 0; pc < p.PC; ++pc) {
                        const float* c0p = strip + (shift + 2 * pc) * kCbufPitch + lane;
                        float mx = -INFINITY;
#pragma unroll
                        for (int a = 0; a < 3; ++a)
#pragma unroll
                            for (int b = 0; b < 3; ++b) mx = fmaxf(mx, c0p[(a * p.TC + b) * kCbufPitch]);
                        float val = fmaxf(mx + bch, 0.f);
                        if (p.round_out) val = to_tf32(val);
                        const int pcg = pc0 + pc;
                        if (row_ok && pcg < g.Wp) {
                            orow[(size_t)pcg * KO] = val;
                            if (pushed) {                       // warp-uniform: only boundary rows take this path
#pragma unroll
                                for (int m = 0; m < B200_MAX_MIRROR; ++m)
                                    if (mrow[m]) mrow[m][(size_t)pcg * KO] = val;
                            }
                        }
                    }
                    __syncwarp();
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar + 8 * buf);
            const int tr0 = pr0, tr1 = min(pr0 + p.PR, g.p1);
            if (g.n_signal > 0 && tile_signals(g, tr0, tr1)) {   // all eight warps' stores, then one signal (CTA-uniform)
                asm volatile("bar.sync 2, 256;" ::: "memory");
                if (warp == 11 && lane == 0) tile_signal(g, n, tr0, tr1);   // quarter 3 has no channels when K_out = 96
            }
            if (++buf == 2) { buf = 0; bphase ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
    if (threadIdx.x == 0) block_done_signal(g);
}


// ------------------------------------------ double-buffered kernel (C % 32 == 0 layers) ----
// Selected with conv2=db (Policy); the fallback of the pair kernels for geometries they refuse.
// tc_block_kernel<false> fills all of TMEM with one 256-pixel tile and its producer warps are its
// epilogue warps, so drain, max-pool and LRN (55 us of conv2's 282 us) never overlap an MMA.  Here a
// tile is 128 pixel slots (one M=128 accumulator of N <= 256 columns), TMEM holds two of them, and
// the roles are split as in tc_block_swapped_kernel:
//   warps 0-7  im2col producers (cp.async 16 B chunks straight into the swizzled A stage, 4 rows each)
//   warp  8    MMA issuer + TMEM owner;   warp 9  TMA issuer (weight k-blocks)
//   warps 10-13 epilogue: TMEM lane quarter warp%4, 32 columns at a time -> cbuf -> pool -> pbuf -> LRN
//               -> store; the accumulator is handed back to the MMA warp BEFORE the LRN pass.
// The epilogue buffers do not alias the stages, so the TMA warp never waits for an epilogue.
constexpr int kDbThreads = 448;
constexpr int kDbEpiWarp0 = 10;
constexpr int kDbTileM = 128;
constexpr int kDbAStageBytes = kDbTileM * 128;            // 16 KB
constexpr int kDbCbufBytes = kDbTileM * kCbufPitch * 4;   // one 32-channel chunk of the tile

struct DbParams {
    TcParams t;                                          // nacc == 1 tiling
    uint32_t cbuf_off, pbuf_off, db_ctl_off;
};

__device__ __forceinline__ void bar_epilogue() { asm volatile("bar.sync 2, 128;" ::: "memory"); }

__global__ void __launch_bounds__(kDbThreads, 1)
tc_block_db_kernel(const __grid_constant__ CUtensorMap tmapB, const __grid_constant__ DbParams dp,
                   const float* __restrict__ bias, float* __restrict__ out) {
    extern __shared__ uint8_t smem_raw[];
    const TcParams& p = dp.t;
    const b200_block_geom& g = p.g;
    const int N = g.K;

    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
    const int NS = p.stages;
    const uint32_t b_off = NS * kDbAStageBytes;
    const uint32_t full_bar = base + dp.db_ctl_off;           // [4]
    const uint32_t empty_bar = full_bar + 32;                 // [4]
    const uint32_t tfull_bar = full_bar + 64;                 // [2] accumulator ready
    const uint32_t tempty_bar = full_bar + 80;                // [2] accumulator drained
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gbase + dp.db_ctl_off + 96);
    const float** rowtab = reinterpret_cast<const float**>(gbase + dp.db_ctl_off + 112);
    int2* ptab = reinterpret_cast<int2*>(gbase + dp.db_ctl_off + 112 + kRowTab * 8);
    float* s_bias = reinterpret_cast<float*>(gbase + dp.db_ctl_off + 112 + kRowTab * 8 + kMaxPooled * 8);
    float* cbuf0 = reinterpret_cast<float*>(gbase + dp.cbuf_off);
    float* pbuf = reinterpret_cast<float*>(gbase + dp.pbuf_off);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; ++s) {
            mbar_init(full_bar + 8 * s, kProducerThreads + 1);
            mbar_init(empty_bar + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(tfull_bar + 8 * b, 1);
            mbar_init(tempty_bar + 8 * b, 4);
        }
        fence_mbar_init();
        block_gate_wait(g);
    }
    for (int i = threadIdx.x; i < N; i += kDbThreads) s_bias[i] = bias[i];
    if (warp == kMmaWarp) tmem_alloc(smem_u32(tmem_slot), p.tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int tiles_per_img = p.tiles_r * p.tiles_c;

    if (warp < kProducerWarps) {
        // =============================== im2col producers (A operand) =====================
        const int tid = threadIdx.x;
        const int j = tid & 7, m_base = tid >> 3;             // rows m_base + 32*i, i < 4
        const uint32_t row_sw = (uint32_t)(m_base & 7);
        const uint32_t dst0 = (uint32_t)(m_base >> 3) * 1024u + row_sw * 128u + (((uint32_t)j ^ row_sw) << 4);
        const int IRT = (p.TR - 1) * g.S + g.F;
        int pij[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int m = m_base + 32 * i;
            const int pi = m / p.TC;
            pij[i] = m < p.TR * p.TC ? (pi << 16) | (m - pi * p.TC) : -1;
        }
        int stage = 0;
        uint32_t phase = 0;
        for (long long w = blockIdx.x; w < p.total_tiles; w += gridDim.x) {
            const int n = (int)(w / tiles_per_img);
            const int rem = (int)(w - (long long)n * tiles_per_img);
            const int tr_i = rem / p.tiles_c, tc_i = rem - tr_i * p.tiles_c;
            const int cr0 = 2 * (g.p0 + tr_i * p.PR), cc0 = 2 * (tc_i * p.PC);
            const int ir0 = cr0 * g.S - g.P;
            if (p.any_flag && tid == 0) {                       // consumer side of the halo flags
                for (int s = 0; s < g.nseg; ++s)
                    if (g.seg[s].flag && ir0 < g.seg[s].hi && ir0 + IRT > g.seg[s].lo)
                        wait_flag_ge(g.seg[s].flag + n, g.seg[s].target * (g.epoch ? *g.epoch : 1), g.status,
                                     g.timeout_ns);
            }
            bar_producers();                                    // previous tile's row-table reads are over
            int rsel[4], cbyte[4];
            uint32_t xmask[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int pi = pij[i] >> 16, pj = pij[i] & 0xFFFF;
                const bool ok = pij[i] >= 0 && (cr0 + pi) < g.Hc && (cc0 + pj) < g.Wc;
                const int x0 = (cc0 + pj) * g.S - g.P;
                rsel[i] = ok ? pi * g.S : IRT;                  // IRT.. = null rows -> zero fill
                cbyte[i] = (x0 * g.C + 4 * j) * 4;
                const int lo = max(0, -x0), hi = min(g.F, g.W - x0);
                xmask[i] = hi > lo ? (((1u << hi) - 1u) & ~((1u << lo) - 1u)) : 0u;
            }
            if (tid < IRT + g.F) {
                const int h = ir0 + tid;
                rowtab[tid] = (tid < IRT && h >= 0 && h < g.H) ? seg_row_ptr(g, n, h) : nullptr;
            }
            bar_producers();
            int fy = 0, fx = 0, cch = 0;
            const char* rp4[4];
            uint32_t vm[4];
            for (int kb = 0; kb < p.nkb; ++kb) {
                if (fx == 0) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        rp4[i] = reinterpret_cast<const char*>(rowtab[rsel[i] + fy]);
                        vm[i] = rp4[i] ? xmask[i] : 0u;
                    }
                }
                mbar_wait(empty_bar + 8 * stage, phase ^ 1);
                const uint32_t a_dst = base + stage * kDbAStageBytes + dst0;
                const int off = (fx * g.C + cch * kBlockK) * 4;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t nb = ((vm[i] >> fx) & 1u) ? 16u : 0u;
                    cp_async16(a_dst + i * 4096u, rp4[i] + (cbyte[i] + off), nb);
                }
                cp_async_mbar_arrive_noinc(full_bar + 8 * stage);
                if (++fx == g.F) { fx = 0; if (++fy == g.F) { fy = 0; ++cch; } }
                if (++stage == NS) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == kMmaWarp) {
        // ======================================= MMA issuer ==============================
        int stage = 0, buf = 0;
        uint32_t phase = 0, bphase = 0;
        for (long long w = blockIdx.x; w < p.total_tiles; w += gridDim.x) {
            mbar_wait(tempty_bar + 8 * buf, bphase ^ 1);        // epilogue has drained this accumulator
            tc_fence_after();
            for (int kb = 0; kb < p.nkb; ++kb) {
                mbar_wait(full_bar + 8 * stage, phase);
                fence_proxy_async_smem();                       // producers' cp.async data -> async-proxy reads
                tc_fence_after();
                if (lane == 0) {
                    const uint64_t da = make_smem_desc(base + stage * kDbAStageBytes);
                    const uint64_t db = make_smem_desc(base + b_off + stage * p.b_stage_bytes);
#pragma unroll
                    for (int ks = 0; ks < kBlockK / 8; ++ks)
                        umma_tf32(tmem_base + (uint32_t)(buf * N), da + 2 * ks, db + 2 * ks, p.idesc,
                                  (kb | ks) != 0 ? 1u : 0u);
                    umma_commit(empty_bar + 8 * stage);
                    if (kb == p.nkb - 1) umma_commit(tfull_bar + 8 * buf);
                }
                __syncwarp();
                if (++stage == NS) { stage = 0; phase ^= 1; }
            }
            if (++buf == 2) { buf = 0; bphase ^= 1; }
        }
    } else if (warp == kTmaWarp) {
        // ======================================= TMA issuer (weights) ====================
        int stage = 0;
        uint32_t phase = 0;
        for (long long w = blockIdx.x; w < p.total_tiles; w += gridDim.x) {
            for (int kb = 0; kb < p.nkb; ++kb) {
                mbar_wait(empty_bar + 8 * stage, phase ^ 1);
                if (lane == 0) {
                    const uint32_t fb = full_bar + 8 * stage;
                    mbar_arrive_expect_tx(fb, (uint32_t)p.b_stage_bytes);
                    tma_load_2d(base + b_off + stage * p.b_stage_bytes, &tmapB, kb * kBlockK, 0, fb);
                }
                __syncwarp();
                if (++stage == NS) { stage = 0; phase ^= 1; }
            }
        }
    } else {
        // ======================================= epilogue warps ==========================
        const int we = warp & 3;                               // TMEM lane quarter this warp may read (warps 10..13)
        epilogue_tables<128>(p, ptab, pbuf, (warp - kDbEpiWarp0) * 32 + lane);
        bar_epilogue();
        int buf = 0;
        uint32_t bphase = 0;
        for (long long w = blockIdx.x; w < p.total_tiles; w += gridDim.x) {
            const int n = (int)(w / tiles_per_img);
            const int rem = (int)(w - (long long)n * tiles_per_img);
            const int tr_i = rem / p.tiles_c, tc_i = rem - tr_i * p.tiles_c;
            const int pr0 = g.p0 + tr_i * p.PR, pc0 = tc_i * p.PC;
            mbar_wait(tfull_bar + 8 * buf, bphase);
            tc_fence_after();
            const uint32_t tb = tempty_bar + 8 * buf;
            epilogue_pool_lrn<2, 4>(p, tmem_base + (uint32_t)(buf * N), cbuf0, kDbCbufBytes / 4, pbuf, ptab, s_bias, out,
                                    n, pr0, pc0, we, lane, true, [&] { if (lane == 0) mbar_arrive(tb); });
            if (++buf == 2) { buf = 0; bphase ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after();
        tmem_dealloc(tmem_base, p.tmem_cols);
    }
    if (threadIdx.x == 0) block_done_signal(g);
}


// ------------------------------------------ CTA-pair kernel (stride-1 layers, C % 32 == 0) ----
// Selected with conv2=pair (Policy); the channel-split pair kernel below is the default.
// tc_block_db_kernel overlaps the epilogue but streams the whole weight k-block (32 KB) into every SM for only
// 128 pixels: 48 KB per 640 MMA cycles is more than an SM can take in from L2, and it runs at ~80 % of the
// MMA rate.  Here two CTAs of a cluster form a tcgen05 pair (cta_group::2): each stages its own 128 pixel rows
// (A) and HALF of the weight rows (B), the leader issues one M=256 x N x K8 MMA per k-step for both, and each
// CTA finds its 128 x N accumulator in its own TMEM, double-buffered as in the db kernel.  The operands are
// fed by TMA only: for a stride-1 convolution the im2col rows of a whole TR x TC tile for one filter tap and
// 32 channels are the box {32, TC, TR, 1} of the NHWC activation at (c, x0+fx-P, y0+fy-P, n), zero-filled
// outside the image — no producer warps, no row table.  Single-GPU geometries only for now (one local row
// segment, no flags, no signals).
//   warp 0: TMA issuer (both CTAs; bytes are counted on the leader's barrier)
//   warp 1: TMEM owner (both CTAs) + MMA issuer (leader only)
//   warps 2-5: epilogue of the CTA's own tile (TMEM lane quarter = warp % 4)
constexpr int kPairThreads = 192;
constexpr int kPairEpiWarp0 = 2;
constexpr int kPairMaxStages = 5;                     // 5 x 32 KB of stages + epilogue buffers still fit

struct PairParams {
    TcParams t;                                          // nacc == 1 tiling (same as the db kernel)
    uint32_t cbuf_off, pbuf_off, pair_ctl_off;
    uint32_t a_box_bytes, b_half_bytes;                  // per CTA and k-block
    uint32_t idesc2;                                     // M = 256
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPairThreads, 1)
tc_block_pair_kernel(const __grid_constant__ CUtensorMap tmapBh, const __grid_constant__ CUtensorMap tmapA,
                     const __grid_constant__ PairParams pp, const float* __restrict__ bias,
                     float* __restrict__ out) {
    extern __shared__ uint8_t smem_raw[];
    const TcParams& p = pp.t;
    const b200_block_geom& g = p.g;
    const int N = g.K;
    const uint32_t rank = cluster_ctarank();                  // 0 = leader (even SM of the pair)

    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
    const int NS = p.stages;
    const uint32_t b_off = NS * kDbAStageBytes;
    const uint32_t full_bar = base + pp.pair_ctl_off;         // [8]  leader's copy is the live one
    const uint32_t empty_bar = full_bar + 64;                 // [8]  per CTA (multicast commit)
    const uint32_t tfull_bar = full_bar + 128;                // [2]  per CTA (multicast commit)
    const uint32_t tempty_bar = full_bar + 144;               // [2]  leader's copy, 8 arrivals (4 warps x 2 CTAs)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gbase + pp.pair_ctl_off + 160);
    int2* ptab = reinterpret_cast<int2*>(gbase + pp.pair_ctl_off + 176);
    float* s_bias = reinterpret_cast<float*>(gbase + pp.pair_ctl_off + 176 + kMaxPooled * 8);
    float* cbuf0 = reinterpret_cast<float*>(gbase + pp.cbuf_off);
    float* pbuf = reinterpret_cast<float*>(gbase + pp.pbuf_off);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; ++s) {
            mbar_init(full_bar + 8 * s, 1);
            mbar_init(empty_bar + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(tfull_bar + 8 * b, 1);
            mbar_init(tempty_bar + 8 * b, 8);
        }
        fence_mbar_init();
        block_gate_wait(g);
    }
    for (int i = threadIdx.x; i < N; i += kPairThreads) s_bias[i] = bias[i];
    if (warp == 1) tmem_alloc2(smem_u32(tmem_slot), p.tmem_cols);
    tc_fence_before();
    cluster_sync_all();                                       // both CTAs' barriers and TMEM exist
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int tiles_per_img = p.tiles_r * p.tiles_c;
    const long long npairs = gridDim.x >> 1, pair_id = blockIdx.x >> 1;

    if (warp == 0) {
        // ======================================= TMA issuer ==============================
        int stage = 0;
        uint32_t phase = 0;
        for (long long pt = pair_id; 2 * pt < p.total_tiles; pt += npairs) {
            long long w = 2 * pt + rank;
            if (w >= p.total_tiles) w = p.total_tiles - 1;     // odd tile count: the partner's tile again
            const int n = (int)(w / tiles_per_img);
            const int rem = (int)(w - (long long)n * tiles_per_img);
            const int tr_i = rem / p.tiles_c, tc_i = rem - tr_i * p.tiles_c;
            const int cr0 = 2 * (g.p0 + tr_i * p.PR), cc0 = 2 * (tc_i * p.PC);
            const int ir0 = cr0 * g.S - g.P;                   // first input row of the tile
            if (p.any_flag) {                                   // pushed halo rows of image n have landed?
                if (lane == 0) {
                    const int IRT = (p.TR - 1) * g.S + g.F;
                    for (int s = 0; s < g.nseg; ++s)
                        if (g.seg[s].flag && ir0 < g.seg[s].hi && ir0 + IRT > g.seg[s].lo)
                            wait_flag_ge(g.seg[s].flag + n, g.seg[s].target * (g.epoch ? *g.epoch : 1), g.status,
                                         g.timeout_ns);
                    fence_proxy_async_global();                 // peer's generic stores -> our TMA reads
                }
                __syncwarp();
            }
            const int y0 = ir0 - g.seg[0].base;                 // slab-relative row of filter row 0
            int fy = 0, fx = 0, cch = 0;
            for (int kb = 0; kb < p.nkb; ++kb) {
                mbar_wait(empty_bar + 8 * stage, phase ^ 1);
                if (lane == 0) {
                    const uint32_t fb = full_bar + 8 * stage;
                    if (rank == 0) mbar_arrive_expect_tx(fb, 2u * (pp.a_box_bytes + pp.b_half_bytes));
                    tma2_load_4d(base + stage * kDbAStageBytes, &tmapA, cch * kBlockK, cc0 * g.S + fx - g.P,
                                 y0 + fy, n, fb);
                    tma2_load_2d(base + b_off + stage * pp.b_half_bytes, &tmapBh, kb * kBlockK,
                                 (int)rank * (N / 2), fb);
                }
                __syncwarp();
                if (++fx == g.F) { fx = 0; if (++fy == g.F) { fy = 0; ++cch; } }
                if (++stage == NS) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ======================================= MMA issuer (leader CTA only) ============
        if (rank == 0) {
            int stage = 0, buf = 0;
            uint32_t phase = 0, bphase = 0;
            for (long long pt = pair_id; 2 * pt < p.total_tiles; pt += npairs) {
                mbar_wait(tempty_bar + 8 * buf, bphase ^ 1);    // both CTAs' epilogues have drained this buffer
                tc_fence_after();
                for (int kb = 0; kb < p.nkb; ++kb) {
                    mbar_wait(full_bar + 8 * stage, phase);
                    tc_fence_after();
                    if (lane == 0) {
                        const uint64_t da = make_smem_desc(base + stage * kDbAStageBytes);
                        const uint64_t db = make_smem_desc(base + b_off + stage * pp.b_half_bytes);
#pragma unroll
                        for (int ks = 0; ks < kBlockK / 8; ++ks)
                            umma2_tf32(tmem_base + (uint32_t)(buf * N), da + 2 * ks, db + 2 * ks, pp.idesc2,
                                       (kb | ks) != 0 ? 1u : 0u);
                        umma2_commit_mc(empty_bar + 8 * stage, 3);       // frees the stage in both CTAs
                        if (kb == p.nkb - 1) umma2_commit_mc(tfull_bar + 8 * buf, 3);
                    }
                    __syncwarp();
                    if (++stage == NS) { stage = 0; phase ^= 1; }
                }
                if (++buf == 2) { buf = 0; bphase ^= 1; }
            }
        }
    } else {
        // ======================================= epilogue warps ==========================
        const int we = warp & 3;                               // TMEM lane quarter this warp may read (warps 2..5)
        epilogue_tables<128>(p, ptab, pbuf, (warp - kPairEpiWarp0) * 32 + lane);
        bar_epilogue();
        int buf = 0;
        uint32_t bphase = 0;
        for (long long pt = pair_id; 2 * pt < p.total_tiles; pt += npairs) {
            const long long w = 2 * pt + rank;
            const bool valid = w < p.total_tiles;               // CTA-uniform
            mbar_wait(tfull_bar + 8 * buf, bphase);
            tc_fence_after();
            const uint32_t tb = tempty_bar + 8 * buf;
            if (!valid) {                                       // duplicate of the partner's tile: drop it
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_leader(tb);
                if (++buf == 2) { buf = 0; bphase ^= 1; }
                continue;
            }
            const int n = (int)(w / tiles_per_img);
            const int rem = (int)(w - (long long)n * tiles_per_img);
            const int tr_i = rem / p.tiles_c, tc_i = rem - tr_i * p.tiles_c;
            const int pr0 = g.p0 + tr_i * p.PR, pc0 = tc_i * p.PC;
            epilogue_pool_lrn<2, 4>(p, tmem_base + (uint32_t)(buf * N), cbuf0, kDbCbufBytes / 4, pbuf, ptab, s_bias, out,
                                    n, pr0, pc0, we, lane, true, [&] { if (lane == 0) mbar_arrive_leader(tb); });
            if (++buf == 2) { buf = 0; bphase ^= 1; }
        }
    }

    tc_fence_before();
    cluster_sync_all();                                       // both CTAs are done with both TMEMs and barriers
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc2(tmem_base, p.tmem_cols);
    }
    if (threadIdx.x == 0) block_done_signal(g);
}

// ------------------------------------------ channel-split CTA-pair kernel (K_out = 256, stride 1) ----
// The pair kernel above gives each CTA its own 128-pixel tile, so a rank's 7 x 13 pooled block (7/13 of an image at 8
// GPUs) needs 5 tiles and 32 images do not fit one round of 74 pairs.  Here the pair shares ONE tile of up to 256 pixel
// slots and splits the CHANNELS instead — the tcgen05 pair issues D[256 channels x 256 pixels] = W[256 x K] * X^T:
//   A operand = weights, M = 256: CTA r stages (and gets the accumulator rows of) channels 128r .. 128r+127
//   B operand = im2col pixels, N = 256: CTA r stages the pixel rows [r*HR, (r+1)*HR) of the tile (HR x TC <= 128 slots,
//               one TMA box per filter tap and 32 input channels, zero fill outside the slab)
//   D in each CTA's TMEM = its 128 channels (lanes) x 256 pixel slots (columns), double buffered (2 x 256 columns).
// A tile is (2PR+1) x (2PC+1) conv pixels = PR x PC pooled outputs (8 x 15 + 7 x 15 pixels for 7 x 7 pooled): the whole
// half-image of a rank is 2 tiles, a whole 27 x 27 image 4.
// Epilogue (4 warps, lane = channel): three conv rows at a time from TMEM through a per-warp strip, max-pool along the
// columns, +bias, ReLU into pbuf[pooled pixel][channel].  LRN's window of 5 channels crosses the CTA boundary at channels
// 127|128: each CTA pushes its two edge channels into the partner's halo buffer (DSMEM stores) and the two epilogues meet
// on a cluster-scope mbarrier once per tile; then LRN and 512 B-per-pixel stores of the CTA's channel half.
// Every tensor map here covers this GPU's memory only (fill_cs refuses anything else).
//   warp 0: TMA (both CTAs)   warp 1: TMEM owner (both), MMA issuer (leader)   warps 2-5: epilogue
constexpr int kCsThreads = 192;
constexpr int kCsEpiWarp0 = 2;
constexpr int kCsMaxStages = 5;
constexpr int kCsPixStageBytes = 128 * 128;           // 128 pixel slots x 32 floats
constexpr int kCsWStageBytes = 128 * 128;             // 128 weight rows x 32 floats
constexpr int kCsStripFloats = 3 * 16 * kCbufPitch;   // three conv rows x 16 columns per epilogue warp
constexpr int kCsPbufPitch = 128 + 4;                 // [2 lower-halo | 128 channels | 2 upper-halo]

struct CsParams {
    TcParams t;                                       // PR, PC, TR = 2PR+1, TC = 2PC+1 (<= 16), tiles, nkb, stages
    int HR;                                           // pixel rows staged per CTA (PR + 1)
    uint32_t strip_off, pbuf_off, xh_off, cs_ctl_off;
    uint32_t pix_bytes;                               // bytes of one pixel box (32 x TC x HR floats)
    uint32_t idesc2;                                  // M = 256, N = 256
    int regpool;                                      // 3x3 pooling in registers (every 16-column row load fits its half)
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kCsThreads, 1)
tc_block_cs_kernel(const __grid_constant__ CUtensorMap tmapW, const __grid_constant__ CUtensorMap tmapX,
                   const __grid_constant__ CsParams cp, const float* __restrict__ bias, float* __restrict__ out) {
    extern __shared__ uint8_t smem_raw[];
    const TcParams& p = cp.t;
    const b200_block_geom& g = p.g;
    const uint32_t rank = cluster_ctarank();                  // 0 = leader; also this CTA's channel half

    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
    const int NS = p.stages;
    const uint32_t w_off = NS * kCsPixStageBytes;             // pixel stages first, then weight stages
    const uint32_t full_bar = base + cp.cs_ctl_off;           // [8]  leader's copy is the live one
    const uint32_t empty_bar = full_bar + 64;                 // [8]  per CTA (multicast commit)
    const uint32_t tfull_bar = full_bar + 128;                // [2]  per CTA (multicast commit)
    const uint32_t tempty_bar = full_bar + 144;               // [2]  leader's copy, 8 arrivals (4 warps x 2 CTAs)
    const uint32_t xbar = full_bar + 160;                     // [2]  per CTA, by tile parity: the partner's halo pushes landed
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gbase + cp.cs_ctl_off + 176);
    float* s_bias = reinterpret_cast<float*>(gbase + cp.cs_ctl_off + 192);       // 128 floats
    float* pbuf = reinterpret_cast<float*>(gbase + cp.pbuf_off);
    float* xh = reinterpret_cast<float*>(gbase + cp.xh_off);  // [2 tile parities][kMaxPooled][2] halo channels from the partner

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; ++s) {
            mbar_init(full_bar + 8 * s, 1);
            mbar_init(empty_bar + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(tfull_bar + 8 * b, 1);
            mbar_init(tempty_bar + 8 * b, 8);
            mbar_init(xbar + 8 * b, 1);
        }
        fence_mbar_init();
    }
    for (int i = threadIdx.x; i < 128; i += kCsThreads) s_bias[i] = bias[128 * rank + i];
    if (warp == 1) tmem_alloc2(smem_u32(tmem_slot), 512);
    // Everything above touches only weights and this CTA's own state, so with a programmatic launch (flags bit 1)
    // it overlaps the tail of the previous kernel in the stream.  From here on its output is read.
    grid_dep_wait();
    if (threadIdx.x == 0) {
        block_gate_wait(g);
        fence_proxy_async_global();                           // the previous grid's generic stores -> our TMA reads
    }
    tc_fence_before();
    cluster_sync_all();                                       // both CTAs' barriers and TMEM exist
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int tiles_per_img = p.tiles_r * p.tiles_c;
    const long long npairs = gridDim.x >> 1, pair_id = blockIdx.x >> 1;

    if (warp == 0) {
        // ======================================= TMA issuer ==============================
        int stage = 0;
        uint32_t phase = 0;
        for (long long w = pair_id; w < p.total_tiles; w += npairs) {
            const int n = (int)(w / tiles_per_img);
            const int rem = (int)(w - (long long)n * tiles_per_img);
            const int tr_i = rem / p.tiles_c, tc_i = rem - tr_i * p.tiles_c;
            const int cr0 = 2 * (g.p0 + tr_i * p.PR), cc0 = 2 * (tc_i * p.PC);
            const int ir0 = cr0 - g.P;                          // first input row of the tile (stride 1)
            if (p.any_flag) {                                   // pushed halo rows of image n have landed?
                if (lane == 0) {
                    const int IRT = p.TR - 1 + g.F;
                    for (int s = 0; s < g.nseg; ++s)
                        if (g.seg[s].flag && ir0 < g.seg[s].hi && ir0 + IRT > g.seg[s].lo)
                            wait_flag_ge(g.seg[s].flag + n, g.seg[s].target * (g.epoch ? *g.epoch : 1), g.status,
                                         g.timeout_ns);
                    fence_proxy_async_global();                 // peer's generic stores -> our TMA reads
                }
                __syncwarp();
            }
            const int y0 = ir0 - g.seg[0].base + (int)rank * cp.HR;   // slab-relative row of this CTA's filter row 0
            int fy = 0, fx = 0, cch = 0;
            for (int kb = 0; kb < p.nkb; ++kb) {
                mbar_wait(empty_bar + 8 * stage, phase ^ 1);
                if (lane == 0) {
                    const uint32_t fb = full_bar + 8 * stage;
                    const bool lpix = !(p.dbg & 64), lwgt = !(p.dbg & 128);     // timing experiments (garbage results)
                    if (rank == 0)
                        mbar_arrive_expect_tx(fb, 2u * ((lpix ? cp.pix_bytes : 0u) + (lwgt ? (uint32_t)kCsWStageBytes : 0u)));
                    if (lpix)
                        tma2_load_4d(base + stage * kCsPixStageBytes, &tmapX, cch * kBlockK, cc0 + fx - g.P, y0 + fy, n, fb);
                    if (lwgt)
                        tma2_load_2d(base + w_off + stage * kCsWStageBytes, &tmapW, kb * kBlockK, (int)rank * 128, fb);
                }
                __syncwarp();
                if (++fx == g.F) { fx = 0; if (++fy == g.F) { fy = 0; ++cch; } }
                if (++stage == NS) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ======================================= MMA issuer (leader CTA only) ============
        if (rank == 0) {
            int stage = 0, buf = 0;
            uint32_t phase = 0, bphase = 0;
            for (long long w = pair_id; w < p.total_tiles; w += npairs) {
                mbar_wait(tempty_bar + 8 * buf, bphase ^ 1);    // both CTAs' epilogues have drained this buffer
                tc_fence_after();
                for (int kb = 0; kb < p.nkb; ++kb) {
                    mbar_wait(full_bar + 8 * stage, phase);
                    tc_fence_after();
                    if (lane == 0) {
                        const uint64_t dw = make_smem_desc(base + w_off + stage * kCsWStageBytes);
                        const uint64_t dx = make_smem_desc(base + stage * kCsPixStageBytes);
#pragma unroll
                        for (int ks = 0; ks < kBlockK / 8; ++ks)
                            if (!(p.dbg & 512))
                                umma2_tf32(tmem_base + (uint32_t)(buf * 256), dw + 2 * ks, dx + 2 * ks, cp.idesc2,
                                           (kb | ks) != 0 ? 1u : 0u);
                        umma2_commit_mc(empty_bar + 8 * stage, 3);       // frees the stage in both CTAs
                        if (kb == p.nkb - 1) umma2_commit_mc(tfull_bar + 8 * buf, 3);
                    }
                    __syncwarp();
                    if (++stage == NS) { stage = 0; phase ^= 1; }
                }
                if (++buf == 2) { buf = 0; bphase ^= 1; }
            }
        }
    } else {
        // ======================================= epilogue warps ==========================
        const int q = warp & 3;                                // TMEM lane quarter = channels 32q..32q+31 of this CTA
        const int te = (warp - kCsEpiWarp0) * 32 + lane;       // thread index inside the epilogue group
        const int cl = q * 32 + lane;                          // channel inside the CTA's half
        const int N = g.K;                                     // 256
        float* strip = reinterpret_cast<float*>(gbase + cp.strip_off) + q * kCsStripFloats;
        const float bch = s_bias[cl];
        const int PP = p.PR * p.PC;
        const bool do_lrn = g.lrn.N_lrn > 0;
        const float al = g.lrn.mode == B200_LRN_ALPHA_OVER_N ? g.lrn.alpha / (float)max(g.lrn.N_lrn, 1) : g.lrn.alpha;
        const float nbeta = -g.lrn.beta;
        // conv row ri of the tile sits in TMEM columns col(ri) .. col(ri)+TC-1 of the buffer; 16 columns are loaded from
        // min(col, 240), so tile column 0 lands in strip column col - start
        auto row_col = [&](int ri) { const int h = ri >= cp.HR ? 1 : 0; return 128 * h + (ri - h * cp.HR) * p.TC; };
        auto row_shift = [&](int ri) { const int c = row_col(ri); return c + 16 > 256 ? c - (256 - 16) : 0; };
        int buf = 0, it = 0;
        uint32_t bphase = 0;
        for (long long w = pair_id; w < p.total_tiles; w += npairs, ++it) {
            const int n = (int)(w / tiles_per_img);
            const int rem = (int)(w - (long long)n * tiles_per_img);
            const int tr_i = rem / p.tiles_c, tc_i = rem - tr_i * p.tiles_c;
            const int pr0 = g.p0 + tr_i * p.PR, pc0 = tc_i * p.PC;
            float* xh_t = xh + (it & 1) * (kMaxPooled * 2);     // this tile's halo buffer (by parity: the partner may
                                                               // already push the next tile's while we read this one)
            mbar_wait(tfull_bar + 8 * buf, bphase);
            tc_fence_after();
            if (cp.regpool && !(p.dbg & 256)) {
                // Lane = channel, TMEM columns = pixels: the 3x3 window is nine registers of one lane.  Three 16-column
                // rows roll through registers (row 2pr+2 of one pooled row is row 2pr of the next); no smem strip.
                uint32_t ra[16], rb[16], rc[16];
                const uint32_t tq = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * 256);
                tmem_ld16(tq + (uint32_t)row_col(0), rc);
                for (int pr = 0; pr < p.PR; ++pr) {
#pragma unroll
                    for (int c = 0; c < 16; ++c) ra[c] = rc[c];
                    tmem_ld16(tq + (uint32_t)row_col(2 * pr + 1), rb);
                    tmem_ld16(tq + (uint32_t)row_col(2 * pr + 2), rc);
                    const int prg = pr0 + pr;
#pragma unroll
                    for (int pc = 0; pc < 7; ++pc) {
                        if (pc < p.PC) {
                            float mx = fmaxf(fmaxf(__uint_as_float(ra[2 * pc]), __uint_as_float(ra[2 * pc + 1])),
                                             __uint_as_float(ra[2 * pc + 2]));
                            mx = fmaxf(mx, fmaxf(fmaxf(__uint_as_float(rb[2 * pc]), __uint_as_float(rb[2 * pc + 1])),
                                                 __uint_as_float(rb[2 * pc + 2])));
                            mx = fmaxf(mx, fmaxf(fmaxf(__uint_as_float(rc[2 * pc]), __uint_as_float(rc[2 * pc + 1])),
                                                 __uint_as_float(rc[2 * pc + 2])));
                            float val = fmaxf(mx + bch, 0.f);
                            const int pp = pr * p.PC + pc;
                            if (do_lrn) {
                                pbuf[pp * kCsPbufPitch + 2 + cl] = val;
                                if (rank == 1 && cl < 2) st_partner_f32(smem_u32(xh_t + pp * 2 + cl), val);
                                if (rank == 0 && cl >= 126) st_partner_f32(smem_u32(xh_t + pp * 2 + (cl - 126)), val);
                            } else {
                                const int pcg = pc0 + pc;
                                if (p.round_out) val = to_tf32(val);
                                if (prg < g.p1 && prg < g.Hp && pcg < g.Wp)
                                    out_ptr_at(out, n, g.out_slab, g.out_base, prg, g.Wp, pcg, N)[128 * rank + cl] = val;
                            }
                        }
                    }
                }
            }
            for (int pr = 0; pr < ((cp.regpool || (p.dbg & 256)) ? 0 : p.PR); ++pr) {
                // conv rows 2pr, 2pr+1, 2pr+2 live in row buffers (row % 3); the first is left over from pr-1
                for (int ri = (pr == 0 ? 0 : 2 * pr + 1); ri <= 2 * pr + 2; ++ri) {
                    const int start = row_col(ri) - row_shift(ri);
                    const int rb = ri % 3;
                    uint32_t r[16];
                    tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * 256 + start), r);
#pragma unroll
                    for (int c = 0; c < 16; ++c) strip[(rb * 16 + c) * kCbufPitch + lane] = __uint_as_float(r[c]);
                }
                __syncwarp();
                const int prg = pr0 + pr;
                const int b0 = (2 * pr) % 3, b1 = (2 * pr + 1) % 3, b2 = (2 * pr + 2) % 3;
                const float* s0 = strip + (b0 * 16 + row_shift(2 * pr)) * kCbufPitch + lane;
                const float* s1 = strip + (b1 * 16 + row_shift(2 * pr + 1)) * kCbufPitch + lane;
                const float* s2 = strip + (b2 * 16 + row_shift(2 * pr + 2)) * kCbufPitch + lane;
                for (int pc = 0; pc < p.PC; ++pc) {
                    const int o = 2 * pc * kCbufPitch;
                    float mx = fmaxf(fmaxf(s0[o], s0[o + kCbufPitch]), s0[o + 2 * kCbufPitch]);
                    mx = fmaxf(mx, fmaxf(fmaxf(s1[o], s1[o + kCbufPitch]), s1[o + 2 * kCbufPitch]));
                    mx = fmaxf(mx, fmaxf(fmaxf(s2[o], s2[o + kCbufPitch]), s2[o + 2 * kCbufPitch]));
                    float val = fmaxf(mx + bch, 0.f);           // relu(max(x)+b) == max(relu(x+b))
                    const int pp = pr * p.PC + pc;
                    if (do_lrn) {
                        pbuf[pp * kCsPbufPitch + 2 + cl] = val;
                        // LRN window halo: channels 128,129 go to CTA 0's upper halo, 126,127 to CTA 1's lower halo
                        if (rank == 1 && cl < 2) st_partner_f32(smem_u32(xh_t + pp * 2 + cl), val);
                        if (rank == 0 && cl >= 126) st_partner_f32(smem_u32(xh_t + pp * 2 + (cl - 126)), val);
                    } else {
                        const int pcg = pc0 + pc;
                        if (p.round_out) val = to_tf32(val);
                        if (prg < g.p1 && prg < g.Hp && pcg < g.Wp)
                            out_ptr_at(out, n, g.out_slab, g.out_base, prg, g.Wp, pcg, N)[128 * rank + cl] = val;
                    }
                }
                __syncwarp();
            }
            // every tcgen05.ld of this warp has completed: hand the accumulator back before the LRN pass
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(tempty_bar + 8 * buf);
            if (do_lrn && !(p.dbg & 256)) {
                bar_epilogue();                                 // pbuf is whole; this CTA's halo pushes are issued
                // two barriers used alternately: the partner can only arrive for tile t+2 after it has seen OUR arrival
                // for t+1, which we issue after this wait — so a barrier never runs two phases ahead of its waiter
                const uint32_t xb = xbar + 8 * (it & 1);
                if (te == 0) mbar_arrive_partner_release(xb);   // ... and now visible to the partner
                mbar_wait_cluster(xb, (uint32_t)((it >> 1) & 1)); // the partner's pushes into OUR halo have landed
                for (int i = te; i < PP * 2; i += 128) {        // halo columns of the pbuf rows (zeros at the layer's edges)
                    const int pp = i >> 1, j = i & 1;
                    pbuf[pp * kCsPbufPitch + j] = rank == 1 ? xh_t[i] : 0.f;
                    pbuf[pp * kCsPbufPitch + 130 + j] = rank == 0 ? xh_t[i] : 0.f;
                }
                bar_epilogue();
                for (int pp = 0; pp < PP; ++pp) {               // thread = channel: 512 B per pixel and CTA
                    const int pi = pp / p.PC, pj = pp - pi * p.PC;
                    const int prg = pr0 + pi, pcg = pc0 + pj;
                    const float* px = pbuf + pp * kCsPbufPitch + te;     // channel te sits at px[2]
                    const float a0 = px[0], a1 = px[1], a2 = px[2], a3 = px[3], a4 = px[4];
                    const float ss = fmaf(a0, a0, fmaf(a1, a1, fmaf(a2, a2, fmaf(a3, a3, a4 * a4))));
                    float val = a2 * exp2f(nbeta * __log2f(g.lrn.k + al * ss));
                    if (p.round_out) val = to_tf32(val);
                    if (prg < g.p1 && prg < g.Hp && pcg < g.Wp)
                        out_ptr_at(out, n, g.out_slab, g.out_base, prg, g.Wp, pcg, N)[128 * rank + te] = val;
                }
                bar_epilogue();                                 // pbuf free for the next tile
            }
            if (++buf == 2) { buf = 0; bphase ^= 1; }
        }
    }

    tc_fence_before();
    cluster_sync_all();                                       // both CTAs are done with both TMEMs and barriers
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc2(tmem_base, 512);
    }
    if (threadIdx.x == 0) block_done_signal(g);
}

// --------------------------------------------------------------------------- host ----
// Which kernel serves a layer is a process-wide policy, decided ONCE (environment at first use, or
// b200_tc_set_policy) and never by a getenv on the launch path:
//   conv2-type layers (C % 32 == 0):  auto = the channel-split CTA-pair kernel where it applies (256 output
//     channels, stride 1, one local slab), else the pixel-split CTA-pair kernel (stride 1, one local slab), else
//     the double-buffered kernel;  "single" / "db" / "pair" / "cs" force one (the pair kernels still fall back
//     on geometries they cannot take)
//   conv1-type layers (C % 32 != 0, K_out <= 128): auto = pixel operand by TMA when the input slab is local and
//     eligible, else cp.async;  "cpasync" forces the LSU loader.  Both loaders use one K order, so they are
//     bit-identical.
struct Policy {
    int conv2 = 0;        // 0 auto, 1 single-buffered, 2 double-buffered, 3 pair, 4 channel-split pair
    int conv1_tma = 1;    // 1 auto (TMA when eligible), 0 cp.async only
    int no_swap = 0;      // conv1-type layers through the generic kernel (experiments)
    int no_direct = 0;    // conv1: shared-memory patch instead of the direct loader (experiments)
    int dbg = 0;          // perf-decomposition knobs of tc_block_kernel (TcParams::dbg)
    int pdl = 1;          // honour flags bit 1 (programmatic dependent launch of the channel-split kernel)
};

int parse_conv2(const char* v) {
    if (!v) return 0;
    if (!strcmp(v, "single")) return 1;
    if (!strcmp(v, "db")) return 2;
    if (!strcmp(v, "pair")) return 3;
    if (!strcmp(v, "cs")) return 4;
    return 0;
}

Policy& policy() {
    static Policy pol = [] {
        Policy q;
        q.conv2 = parse_conv2(getenv("B200_TC_CONV2"));
        if (const char* e = getenv("B200_TC_CONV1")) q.conv1_tma = strcmp(e, "cpasync") != 0;
        q.no_swap = getenv("B200_TC_NO_SWAP") != nullptr;
        q.no_direct = getenv("B200_TC_NO_DIRECT") != nullptr;
        if (const char* e = getenv("B200_TC_DBG")) q.dbg = atoi(e);
        if (const char* e = getenv("B200_TC_PDL")) q.pdl = atoi(e);
        return q;
    }();
    return pol;
}

struct TcConv {
    int K, C, F, Kpad, device, patch_mode, FCpad;
    int sms;                                             // SMs of `device`
    float* d_w;
    CUtensorMap tmap;
    // pixel-TMA mode (draft): weights in the K order (fy | run floats 0..31), then tail blocks (8 filter rows x
    // 4 floats from run offset 32); 0 k-blocks when the shape does not qualify
    int nkb_tma;
    float* d_w_tma;
    CUtensorMap tmap_tma;
    // CTA-pair kernel (draft): the same vec-mode weights with a TMA box of K/2 rows (one half per CTA)
    int half_ok;
    CUtensorMap tmap_half;
};

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            return nullptr;
        fn = (EncodeTiledFn)f;
    }
    return fn;
}

inline float host_tf32(float x) {                     // round to nearest, ties away (cvt.rna)
    uint32_t u;
    memcpy(&u, &x, 4);
    if ((u & 0x7F800000u) != 0x7F800000u) u = (u + 0x1000u) & ~0x1FFFu;
    memcpy(&x, &u, 4);
    return x;
}

inline int cdiv(int a, int b) { return (a + b - 1) / b; }
inline bool is_patch_mode(int C) { return C % kBlockK != 0; }
inline int fcpad(int F, int C) { return (F * C + 3) & ~3; }
inline int num_kblocks(int F, int C) {
    return is_patch_mode(C) ? cdiv(F * fcpad(F, C), kBlockK) : (C / kBlockK) * F * F;
}

constexpr int kMaxPatchBytes = 64 * 1024;

// Tile planner: PR x PC pooled outputs per tile, (2PR+1)(2PC+1) <= 128*nacc pixel slots.
// Cost = waves x cycles-per-tile, with cycles calibrated on B200 (profiles/tc_experiments_v4.log):
// the tensor pipe runs ~1.3x above its floor, a tile carries ~14k cycles of fixed cost (setup, pipeline
// fill/drain) plus an epilogue that grows with accumulators x channel chunks (+LRN), and the grid is
// 148 persistent CTAs, so 1.7 waves cost 2.  Deterministic: peers call it to predict tile counts.
constexpr int kPlanSms = 148;

// small-K_out layers whose B rows can be cp.async'd straight from the activation (see SwParams::direct)
// A filter-row run = fcpad(F,C) floats starting at column x*S*C of an input row; it is made of aligned
// 16 B chunks when S*C and the row pitch are multiples of 4 floats, and it stays inside the row when the
// pitch covers the last pixel's run (the floats past W*C only ever meet zero weights).
inline bool direct_eligible(const b200_block_geom* g) {
    const int pitch = g->in_pitch ? g->in_pitch : g->W * g->C;
    return is_patch_mode(g->C) && g->K <= 128 && g->lrn.N_lrn == 0 && g->P == 0 && (g->S * g->C) % 4 == 0 &&
           pitch % 4 == 0 && pitch >= (g->Wc - 1) * g->S * g->C + fcpad(g->F, g->C) &&
           !policy().no_direct;
}

bool plan_tiles(const b200_block_geom* g, TcParams* p, int max_nacc = 2) {
    const int rows = g->p1 - g->p0;
    const bool patch = is_patch_mode(g->C) && !direct_eligible(g);
    const int nkb = num_kblocks(g->F, g->C);
    const double mma_kb = 4.0 * (128.0 * g->K / 256.0) * 1.3;         // cycles per k-block per accumulator
    const double prod_kb = patch ? 450.0 : 0.0;                       // smem->smem im2col bound (conv1)
    double best = -1;
    for (int nacc = 1; nacc <= max_nacc; ++nacc)
        for (int PR = 1; PR <= rows && PR <= 63; ++PR)
            for (int PC = 1; PC <= g->Wp; ++PC) {
                const int TR = 2 * PR + 1, TC = 2 * PC + 1;
                const int slots = TR * TC;
                if (slots > 128 * nacc) break;
                if (PR * PC > kMaxPooled) break;
                if ((TR - 1) * g->S + 2 * g->F > kRowTab) break;
                if (patch) {
                    const int prow = (TR - 1) * g->S + g->F + 1;
                    const int pitch = ((TC - 1) * g->S * g->C + fcpad(g->F, g->C) + 3) & ~3;
                    if (prow * pitch * 4 > kMaxPatchBytes || prow > kRowTab) break;
                }
                const long tiles = (long)cdiv(rows, PR) * cdiv(g->Wp, PC);
                const double per_kb = nacc * (mma_kb > prod_kb ? mma_kb : prod_kb);
                const double tile_cycles = nkb * per_kb + 14000.0 + nacc * (g->K / 32) * 900.0 +
                                           (g->lrn.N_lrn > 0 ? 20000.0 : 0.0);
                const long waves = (tiles * (long)g->N + kPlanSms - 1) / kPlanSms;
                // conv1 kernel: tiles at most 15 conv columns wide pool in registers (no smem strips, deeper ring)
                const double wide = (is_patch_mode(g->C) && g->K <= 128 && g->lrn.N_lrn == 0 && TC > 15) ? 0.5 : 0.0;
                const double cost = waves * tile_cycles + slots * 1e-3 + wide;
                if (best < 0 || cost < best) {
                    best = cost;
                    p->PR = PR; p->PC = PC; p->TR = TR; p->TC = TC; p->nacc = nacc;
                    p->tiles_r = cdiv(rows, PR); p->tiles_c = cdiv(g->Wp, PC);
                }
            }
    return best >= 0;
}

int fill_params(const b200_block_geom* g, TcParams* p, int flags, int max_nacc = 2) {
    if (g->K % 32 != 0 || g->K > 256 || g->K < 32) return B200_ERR_UNSUPPORTED;
    const bool patch = is_patch_mode(g->C);
    if (patch && (g->S * g->C) % 4 != 0) return B200_ERR_UNSUPPORTED;   // LDS.128 alignment
    memset(p, 0, sizeof(*p));
    p->g = *g;
    if (!plan_tiles(g, p, max_nacc)) return B200_ERR_ARG;
    p->nkb = num_kblocks(g->F, g->C);
    p->FCpad = fcpad(g->F, g->C);
    p->round_out = flags & 1;
    p->dbg = policy().dbg;
    p->b_stage_bytes = g->K * 128;
    size_t patch_bytes = 0;
    if (patch) {
        p->patch_rows = (p->TR - 1) * g->S + g->F + 1;
        p->patch_pitch = ((p->TC - 1) * g->S * g->C + p->FCpad + 3) & ~3;
        patch_bytes = (size_t)p->patch_rows * p->patch_pitch * 4;
    }
    const size_t ctl_bytes = 16 * 4 + 32 + kRowTab * 8 + kMaxPooled * 8 + (size_t)g->K * 4 + 64;
    const size_t epi = 2 * (size_t)kCbufBytes + (g->lrn.N_lrn > 0 ? (size_t)kMaxPooled * (g->K + 4) * 4 : 0);
    const int stage_bytes = kAStageBytes + p->b_stage_bytes;
    const long avail = 226 * 1024 - 1024 - (long)patch_bytes - (long)ctl_bytes;
    int stages = (int)(avail / stage_bytes);
    if (stages > 4) stages = 4;
    if (stages < 2 || (size_t)stages * stage_bytes < epi) return B200_ERR_UNSUPPORTED;
    p->stages = stages;
    p->patch_off = (uint32_t)((size_t)stages * stage_bytes);
    p->ctl_off = (uint32_t)((p->patch_off + patch_bytes + 15) & ~(size_t)15);
    int cols = 32;
    while (cols < 2 * g->K) cols *= 2;                 // two accumulators of N columns
    p->tmem_cols = cols;
    for (int s = 0; s < g->nseg; ++s)
        if (g->seg[s].flag) p->any_flag = 1;
    // instruction descriptor: D=f32, A=B=tf32, both K-major, M=128, N=K
    p->idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(g->K >> 3) << 17) | ((128u >> 4) << 24);
    p->total_tiles = (long long)g->N * p->tiles_r * p->tiles_c;
    return B200_OK;
}

// swapped kernel: layout [pixel stages | weight stages | patch | 4 epilogue strips | ctl]
size_t smem_bytes_swapped(const SwParams& sp) { return 1024 + (size_t)sp.sw_ctl_off + kSwCtlBytes + 16; }

// smem layout of the conv1 kernel for a given ring depth: [weight slots | pixel stages | patch | strips | ctl]
int layout_swapped(SwParams* sp, int nstages) {
    TcParams& p = sp->t;
    if (nstages > kSwMaxStages) return B200_ERR_UNSUPPORTED;
    sp->nstages = nstages;
    sp->w_stage_bytes = (uint32_t)((p.g.K * 128 + 1023) & ~1023);
    const size_t stages = (size_t)nstages * (kAStageBytes + sp->w_stage_bytes);
    const size_t patch_bytes = sp->direct ? 0 : (size_t)p.patch_rows * p.patch_pitch * 4;
    p.patch_off = (uint32_t)stages;
    sp->strip_off = (uint32_t)((stages + patch_bytes + 15) & ~(size_t)15);
    sp->strip_bytes = sp->regpool ? 0u : (uint32_t)(32 * sp->nload * kCbufPitch * 4);
    sp->sw_ctl_off = (uint32_t)((sp->strip_off + kSwEpiWarps * (size_t)sp->strip_bytes + 15) & ~(size_t)15);
    // the MMA reads 128 weight rows per slot: the last slot's rows K_out..127 must still be ours (pixel stages follow)
    return smem_bytes_swapped(*sp) > 227 * 1024 ? B200_ERR_UNSUPPORTED : B200_OK;
}

int fill_swapped(const b200_block_geom* g, SwParams* sp, int flags) {
    int rc = fill_params(g, &sp->t, flags);
    if (rc != B200_OK) return rc;
    TcParams& p = sp->t;
    if (!is_patch_mode(g->C) || g->K > 128 || g->lrn.N_lrn > 0) return B200_ERR_UNSUPPORTED;
    sp->nload = cdiv(3 * p.TC, 32);
    if (32 * sp->nload > 128 * p.nacc) return B200_ERR_UNSUPPORTED;
    // direct mode: every 16 B chunk of a B row is an aligned piece of the activation (no padding, channel
    // and row sizes multiples of 4 floats): cp.async from global, no smem patch, no smem->smem im2col
    sp->direct = (direct_eligible(g) && (p.TR - 1) * g->S + g->F <= kRowTab) ? 1 : 0;
    sp->cpr = fcpad(g->F, g->C) / 4;
    // register pooling: a pooled row is three 16-column TMEM loads that stay inside the accumulator
    sp->regpool = (p.PC <= 7 && (p.TR - 1) * p.TC + 16 <= 128 * p.nacc && !(policy().dbg & 2048)) ? 1 : 0;
    // D = channels x pixels: M = 128 (K_out rows padded), N = pixel slots
    sp->sw_idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)((128 * p.nacc) >> 3) << 17) | ((128u >> 4) << 24);
    return layout_swapped(sp, sp->direct ? 3 : 2);
}

// double-buffered kernel: layout [A stages | B stages | 2 cbuf | pbuf | ctl]
// conv2-type layers run 128-pixel tiles (db / pair kernels) unless the single-buffered kernel is forced
inline bool db_requested(const b200_block_geom* g) { return !is_patch_mode(g->C) && policy().conv2 != 1; }   // incl. cs fallback

int fill_db(const b200_block_geom* g, DbParams* dp, int flags) {
    if (is_patch_mode(g->C)) return B200_ERR_UNSUPPORTED;
    int rc = fill_params(g, &dp->t, flags, /*max_nacc=*/1);
    if (rc != B200_OK) return rc;
    TcParams& p = dp->t;
    const size_t pbuf_bytes = g->lrn.N_lrn > 0 ? (size_t)p.PR * p.PC * (g->K + 4) * 4 : 0;
    const size_t epi = 2 * (size_t)kDbCbufBytes + ((pbuf_bytes + 15) & ~(size_t)15);
    const size_t ctl = 112 + kRowTab * 8 + kMaxPooled * 8 + (size_t)g->K * 4 + 64;
    const size_t stage_bytes = (size_t)kDbAStageBytes + p.b_stage_bytes;
    const long avail = 226 * 1024 - 1024 - (long)epi - (long)ctl;
    int stages = (int)(avail / (long)stage_bytes);
    if (stages > 4) stages = 4;
    if (stages < 2) return B200_ERR_UNSUPPORTED;
    p.stages = stages;
    dp->cbuf_off = (uint32_t)(stages * stage_bytes);
    dp->pbuf_off = dp->cbuf_off + 2 * kDbCbufBytes;
    dp->db_ctl_off = (uint32_t)(stages * stage_bytes + epi);
    int cols = 32;
    while (cols < 2 * g->K) cols *= 2;                 // two accumulators of N columns
    p.tmem_cols = cols;
    return B200_OK;
}

size_t smem_bytes_db(const DbParams& dp) {
    return 1024 + (size_t)dp.db_ctl_off + 112 + kRowTab * 8 + kMaxPooled * 8 + (size_t)dp.t.g.K * 4 + 64;
}

// pixel-TMA mode: turn a direct-mode plan into a TMA-fed one and encode the S row-phase maps.  Returns false
// (leave sp untouched) when the shape, the geometry or the driver does not take it.
bool try_pixel_tma(const TcConv* t, const b200_block_geom* g, SwParams* sp, PixMaps* pm) {
    if (!policy().conv1_tma || !sp->direct || t->nkb_tma == 0 || g->nseg != 1 || g->seg[0].flag) return false;
    {   // TMA only over memory of this GPU (an input that lives on a peer is pulled by the cp.async loader)
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, g->seg[0].ptr) != cudaSuccess || at.type != cudaMemoryTypeDevice ||
            at.device != t->device) {
            (void)cudaGetLastError();
            return false;
        }
    }
    if (g->S < 1 || g->S > 4 || g->seg[0].base % g->S != 0 || g->P != 0) return false;
    if (g->F < 4 || t->nkb_tma - g->F > 2) return false;             // tail handshake: two barriers, ring no deeper than F
    const TcParams& p = sp->t;
    if (p.TC > 256 || p.TR > 256) return false;
    EncodeTiledFn enc = get_encode();
    if (!enc) return false;
    const int pitch = g->in_pitch ? g->in_pitch : g->W * g->C;
    const int slab = g->seg[0].slab;
    if ((pitch * 4) % 16 != 0 || ((long long)slab * pitch * 4) % 16 != 0) return false;
    PixMaps maps;
    for (int r = 0; r < g->S; ++r) {
        const int rows = slab > r ? (slab - r + g->S - 1) / g->S : 0;
        if (rows == 0) return false;
        cuuint64_t dims[4] = {(cuuint64_t)kBlockK, (cuuint64_t)g->Wc, (cuuint64_t)rows, (cuuint64_t)g->N};
        cuuint64_t strides[3] = {(cuuint64_t)g->S * g->C * 4, (cuuint64_t)g->S * pitch * 4,
                                 (cuuint64_t)slab * pitch * 4};
        cuuint32_t box[4] = {(cuuint32_t)kBlockK, (cuuint32_t)p.TC, (cuuint32_t)p.TR, 1};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        void* base = (void*)(g->seg[0].ptr + (size_t)r * pitch);
        if (enc(&maps.m[r], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, base, dims, strides, box, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return false;                                            // e.g. overlapping strides refused
    }
    for (int r = g->S; r < 4; ++r) maps.m[r] = maps.m[0];
    *pm = maps;
    sp->pix_bytes = (uint32_t)(kBlockK * p.TC * p.TR * 4);
    if (!(policy().dbg & 1024)) {                     // as deep a ring as fits (the strips' smem, when pooling in registers)
        int ns = sp->nstages;
        while (ns < kSwMaxStages && ns < g->F && layout_swapped(sp, ns + 1) == B200_OK) ++ns;
        layout_swapped(sp, ns);
    }
    if (getenv("B200_TC_VERBOSE"))
        fprintf(stderr, "[b200 tc] conv1: pixel operand by TMA, tile %dx%d conv pixels, %d stages, pooling in %s\n",
                p.TR, p.TC, sp->nstages, sp->regpool ? "registers" : "smem strips");
    return true;
}

// CTA-pair kernel: layout [A stages | B half stages | 2 cbuf | pbuf | ctl]
inline bool pair_requested(const b200_block_geom* g) {
    return !is_patch_mode(g->C) && (policy().conv2 == 0 || policy().conv2 == 3);
}

// true when every row segment is a window of ONE local slab (same pointer, base and height): the halo rows were
// pushed into this rank's own copy, so a single tensor map covers everything conv reads
bool one_local_slab(const TcConv* t, const b200_block_geom* g) {
    for (int s = 1; s < g->nseg; ++s)
        if (g->seg[s].ptr != g->seg[0].ptr || g->seg[s].base != g->seg[0].base || g->seg[s].slab != g->seg[0].slab)
            return false;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, g->seg[0].ptr) != cudaSuccess || at.type != cudaMemoryTypeDevice ||
        at.device != t->device) {
        (void)cudaGetLastError();
        return false;
    }
    return true;
}

int fill_pair(const TcConv* t, const b200_block_geom* g, PairParams* pp, CUtensorMap* tmapA, int flags) {
    if (is_patch_mode(g->C) || !t->half_ok || g->S != 1 || t->sms < 2) return B200_ERR_UNSUPPORTED;
    if (!one_local_slab(t, g)) return B200_ERR_UNSUPPORTED;
    int rc = fill_params(g, &pp->t, flags, /*max_nacc=*/1);
    if (rc != B200_OK) return rc;
    TcParams& p = pp->t;
    if (p.TC > 256 || p.TR > 256) return B200_ERR_UNSUPPORTED;
    pp->b_half_bytes = (uint32_t)(g->K / 2) * 128u;
    pp->a_box_bytes = (uint32_t)(kBlockK * p.TC * p.TR * 4);
    const size_t pbuf_bytes = g->lrn.N_lrn > 0 ? (size_t)p.PR * p.PC * (g->K + 4) * 4 : 0;
    const size_t epi = 2 * (size_t)kDbCbufBytes + ((pbuf_bytes + 15) & ~(size_t)15);
    const size_t ctl = 176 + kMaxPooled * 8 + (size_t)g->K * 4 + 64;
    const size_t stage_bytes = (size_t)kDbAStageBytes + pp->b_half_bytes;
    const long avail = 226 * 1024 - 1024 - (long)epi - (long)ctl;
    int stages = (int)(avail / (long)stage_bytes);
    if (stages > kPairMaxStages) stages = kPairMaxStages;
    if (stages < 2) return B200_ERR_UNSUPPORTED;
    p.stages = stages;
    pp->cbuf_off = (uint32_t)(stages * stage_bytes);
    pp->pbuf_off = pp->cbuf_off + 2 * kDbCbufBytes;
    pp->pair_ctl_off = (uint32_t)(stages * stage_bytes + epi);
    int cols = 32;
    while (cols < 2 * g->K) cols *= 2;
    p.tmem_cols = cols;
    pp->idesc2 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(g->K >> 3) << 17) | ((256u >> 4) << 24);
    // A operand: the NHWC slab itself (rows [base, base+slab) of every image), box = one filter tap x 32 channels
    // of a whole TR x TC tile; rows and columns outside the slab read as zero, which is the conv's padding at the
    // image border and never reaches a valid output elsewhere (the slab holds every row this rank's pixels need)
    EncodeTiledFn enc = get_encode();
    if (!enc) return B200_ERR_UNSUPPORTED;
    const int pitch = g->in_pitch ? g->in_pitch : g->W * g->C;
    const int slab = g->seg[0].slab;
    if (slab <= 0 || (pitch * 4) % 16 != 0) return B200_ERR_UNSUPPORTED;
    cuuint64_t dims[4] = {(cuuint64_t)g->C, (cuuint64_t)g->W, (cuuint64_t)slab, (cuuint64_t)g->N};
    cuuint64_t strides[3] = {(cuuint64_t)g->C * 4, (cuuint64_t)pitch * 4, (cuuint64_t)slab * pitch * 4};
    cuuint32_t box[4] = {(cuuint32_t)kBlockK, (cuuint32_t)p.TC, (cuuint32_t)p.TR, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    if (enc(tmapA, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (void*)g->seg[0].ptr, dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return B200_ERR_UNSUPPORTED;
    return B200_OK;
}

// channel-split pair kernel: layout [pixel stages | weight stages | 4 strips | pbuf | halo exchange | ctl]
inline bool cs_requested(const b200_block_geom* g) {
    return !is_patch_mode(g->C) && (policy().conv2 == 0 || policy().conv2 == 4);
}

int fill_cs(const TcConv* t, const b200_block_geom* g, CsParams* cp, CUtensorMap* tmapX, int flags) {
    if (is_patch_mode(g->C) || !t->half_ok || g->S != 1 || g->K != 256 || t->sms < 2) return B200_ERR_UNSUPPORTED;
    if (g->lrn.N_lrn != 0 && g->lrn.N_lrn != 5) return B200_ERR_UNSUPPORTED;   // halo of two channels per side
    if (g->n_signal != 0 || g->n_mirror != 0) return B200_ERR_UNSUPPORTED;     // a tile's stores come from two CTAs
    if (!one_local_slab(t, g)) return B200_ERR_UNSUPPORTED;                    // TMA reads this GPU's memory only
    memset(cp, 0, sizeof(*cp));
    TcParams& p = cp->t;
    p.g = *g;
    // tile plan: PR x PC pooled per tile, each CTA stages HR = PR+1 pixel rows of TC = 2PC+1 <= 16 columns
    const int rows = g->p1 - g->p0, pairs = t->sms / 2;
    long best_waves = -1, best_tiles = 0;
    int best_slots = 0;
    for (int PR = 1; PR <= rows && PR <= 31; ++PR)
        for (int PC = 1; PC <= 7 && PC <= g->Wp; ++PC) {
            const int HR = PR + 1, TC = 2 * PC + 1;
            if (HR * TC > 128 || PR * PC > kMaxPooled) continue;
            const long tiles = (long)cdiv(rows, PR) * cdiv(g->Wp, PC);
            const long waves = (tiles * g->N + pairs - 1) / pairs;
            const int slots = 2 * HR * TC;
            if (best_waves < 0 || waves < best_waves || (waves == best_waves && (tiles < best_tiles ||
                (tiles == best_tiles && slots < best_slots)))) {
                best_waves = waves; best_tiles = tiles; best_slots = slots;
                p.PR = PR; p.PC = PC; p.TR = 2 * PR + 1; p.TC = TC; p.nacc = 1;
                p.tiles_r = cdiv(rows, PR); p.tiles_c = cdiv(g->Wp, PC);
                cp->HR = HR;
            }
        }
    if (best_waves < 0) return B200_ERR_UNSUPPORTED;
    p.nkb = num_kblocks(g->F, g->C);
    p.round_out = flags & 1;
    p.dbg = policy().dbg;
    p.total_tiles = (long long)g->N * p.tiles_r * p.tiles_c;
    for (int s = 0; s < g->nseg; ++s)
        if (g->seg[s].flag) p.any_flag = 1;
    cp->pix_bytes = (uint32_t)(kBlockK * p.TC * cp->HR * 4);
    cp->regpool = (p.PC <= 7 && (cp->HR - 1) * p.TC + 16 <= 128 && !(policy().dbg & 2048)) ? 1 : 0;
    const size_t strips = 4 * (size_t)kCsStripFloats * 4;
    const size_t pbuf_bytes = (size_t)kMaxPooled * kCsPbufPitch * 4;
    const size_t xh_bytes = 2 * (size_t)kMaxPooled * 2 * 4;
    const size_t ctl = 192 + 128 * 4 + 64;
    const long avail = 226 * 1024 - 1024 - (long)(strips + pbuf_bytes + xh_bytes + ctl);
    int stages = (int)(avail / (kCsPixStageBytes + kCsWStageBytes));
    if (stages > kCsMaxStages) stages = kCsMaxStages;
    if (stages < 2) return B200_ERR_UNSUPPORTED;
    p.stages = stages;
    cp->strip_off = (uint32_t)(stages * (kCsPixStageBytes + kCsWStageBytes));
    cp->pbuf_off = cp->strip_off + (uint32_t)strips;
    cp->xh_off = cp->pbuf_off + (uint32_t)pbuf_bytes;
    cp->cs_ctl_off = cp->xh_off + (uint32_t)xh_bytes;
    p.tmem_cols = 512;
    cp->idesc2 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(256 >> 3) << 17) | ((256u >> 4) << 24);
    EncodeTiledFn enc = get_encode();
    if (!enc) return B200_ERR_UNSUPPORTED;
    const int pitch = g->in_pitch ? g->in_pitch : g->W * g->C;
    const int slab = g->seg[0].slab;
    if (slab <= 0 || (pitch * 4) % 16 != 0) return B200_ERR_UNSUPPORTED;
    cuuint64_t dims[4] = {(cuuint64_t)g->C, (cuuint64_t)g->W, (cuuint64_t)slab, (cuuint64_t)g->N};
    cuuint64_t strides[3] = {(cuuint64_t)g->C * 4, (cuuint64_t)pitch * 4, (cuuint64_t)slab * pitch * 4};
    cuuint32_t box[4] = {(cuuint32_t)kBlockK, (cuuint32_t)p.TC, (cuuint32_t)cp->HR, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    if (enc(tmapX, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (void*)g->seg[0].ptr, dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return B200_ERR_UNSUPPORTED;
    return B200_OK;
}

size_t smem_bytes_cs(const CsParams& cp) { return 1024 + (size_t)cp.cs_ctl_off + 192 + 128 * 4 + 64; }

size_t smem_bytes_pair(const PairParams& pp) {
    return 1024 + (size_t)pp.pair_ctl_off + 176 + kMaxPooled * 8 + (size_t)pp.t.g.K * 4 + 64;
}



size_t smem_bytes(const TcParams& p) {
    return 1024 + (size_t)p.ctl_off + 16 * p.stages + 32 + kRowTab * 8 + kMaxPooled * 8 +
           (size_t)p.g.K * 4 + 64;
}

}  // namespace

extern "C" int b200_tc_conv_create(const float* w_kcff, int K, int C, int F, void** handle) {
    if (!w_kcff || !handle || K <= 0 || C <= 0 || F <= 0) return B200_ERR_ARG;
    EncodeTiledFn enc = get_encode();
    if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return B200_ERR_CUDA; }
    TcConv* t = new TcConv();
    t->K = K; t->C = C; t->F = F;
    t->patch_mode = is_patch_mode(C) ? 1 : 0;
    t->FCpad = fcpad(F, C);
    t->Kpad = num_kblocks(F, C) * kBlockK;
    B200_CUDA_TRY(cudaGetDevice(&t->device));
    B200_CUDA_TRY(cudaDeviceGetAttribute(&t->sms, cudaDevAttrMultiProcessorCount, t->device));
    // [K][C][F][F] -> K-major rows [K][Kpad], TF32-rounded, zero padded; k order per mode (see kernel)
    std::vector<float> wk((size_t)K * t->Kpad, 0.f);
    for (int k = 0; k < K; ++k)
        for (int c = 0; c < C; ++c)
            for (int fy = 0; fy < F; ++fy)
                for (int fx = 0; fx < F; ++fx) {
                    const int kk = t->patch_mode
                                       ? fy * t->FCpad + fx * C + c
                                       : ((c / kBlockK) * F * F + fy * F + fx) * kBlockK + c % kBlockK;
                    wk[(size_t)k * t->Kpad + kk] =
                        host_tf32(w_kcff[(((size_t)k * C + c) * F + fy) * F + fx]);
                }
    B200_CUDA_TRY(cudaMalloc(&t->d_w, wk.size() * sizeof(float)));
    B200_CUDA_TRY(cudaMemcpy(t->d_w, wk.data(), wk.size() * sizeof(float), cudaMemcpyHostToDevice));
    cuuint64_t dims[2] = {(cuuint64_t)t->Kpad, (cuuint64_t)K};
    cuuint64_t strides[1] = {(cuuint64_t)t->Kpad * sizeof(float)};
    cuuint32_t box[2] = {(cuuint32_t)kBlockK, (cuuint32_t)K};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&t->tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, t->d_w, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed: %d (K=%d Kpad=%d)", (int)r, K, t->Kpad);
        cudaFree(t->d_w);
        delete t;
        return B200_ERR_CUDA;
    }
    t->half_ok = 0;
    if (!t->patch_mode && K % 32 == 0 && K <= 256) {                 // CTA-pair kernel: each CTA stages K/2 weight rows
        cuuint32_t hbox[2] = {(cuuint32_t)kBlockK, (cuuint32_t)(K / 2)};
        if (enc(&t->tmap_half, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, t->d_w, dims, strides, hbox, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS)
            t->half_ok = 1;
    }
    // pixel-TMA K order: only for runs of 33..36 floats (one 16 B tail chunk per filter row), K_out <= 128
    t->nkb_tma = 0;
    t->d_w_tma = nullptr;
    if (t->patch_mode && K <= 128 && F * C > kBlockK && t->FCpad <= kBlockK + 4) {
        const int nkb = F + cdiv(F, 8);
        const int kp = nkb * kBlockK;
        std::vector<float> wt((size_t)K * kp, 0.f);
        for (int k = 0; k < K; ++k)
            for (int c = 0; c < C; ++c)
                for (int fy = 0; fy < F; ++fy)
                    for (int fx = 0; fx < F; ++fx) {
                        const int idx = fx * C + c;                   // float inside the filter-row run
                        const int kk = idx < kBlockK ? fy * kBlockK + idx
                                                     : (F + fy / 8) * kBlockK + (fy % 8) * 4 + (idx - kBlockK);
                        wt[(size_t)k * kp + kk] = host_tf32(w_kcff[(((size_t)k * C + c) * F + fy) * F + fx]);
                    }
        if (cudaMalloc(&t->d_w_tma, wt.size() * sizeof(float)) == cudaSuccess &&
            cudaMemcpy(t->d_w_tma, wt.data(), wt.size() * sizeof(float), cudaMemcpyHostToDevice) == cudaSuccess) {
            cuuint64_t d2[2] = {(cuuint64_t)kp, (cuuint64_t)K};
            cuuint64_t s2[1] = {(cuuint64_t)kp * sizeof(float)};
            if (enc(&t->tmap_tma, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, t->d_w_tma, d2, s2, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS)
                t->nkb_tma = nkb;
        }
        (void)cudaGetLastError();                                     // optional path: never poison the error state
    }
    *handle = t;
    return B200_OK;
}

extern "C" int b200_tc_conv_destroy(void* handle) {
    TcConv* t = (TcConv*)handle;
    if (t) {
        cudaFree(t->d_w);
        if (t->d_w_tma) cudaFree(t->d_w_tma);
        delete t;
    }
    return B200_OK;
}

// key: "conv2" = auto|single|db|pair, "conv1" = auto|cpasync, "no_swap" / "no_direct" / "dbg" = integer
extern "C" int b200_tc_set_policy(const char* key, const char* value) {
    if (!key || !value) return B200_ERR_ARG;
    Policy& q = policy();
    if (!strcmp(key, "conv2")) q.conv2 = parse_conv2(value);
    else if (!strcmp(key, "conv1")) q.conv1_tma = strcmp(value, "cpasync") != 0;
    else if (!strcmp(key, "no_swap")) q.no_swap = atoi(value);
    else if (!strcmp(key, "no_direct")) q.no_direct = atoi(value);
    else if (!strcmp(key, "dbg")) q.dbg = atoi(value);
    else if (!strcmp(key, "pdl")) q.pdl = atoi(value);
    else return B200_ERR_ARG;
    return B200_OK;
}

// tiles per image that produce at least one pooled row of [lo,hi): the arrival-counter target of a reader that
// only waits for those rows (b200_block_geom::signal_lo / signal_hi)
extern "C" int b200_tc_block_tiles_rows(const b200_block_geom* g, int lo, int hi) {
    if (!g) return B200_ERR_ARG;
    if (g->p1 <= g->p0) return 0;
    TcParams p;
    bool have = false;
    if (db_requested(g) || pair_requested(g)) {
        DbParams dp;
        if (fill_db(g, &dp, 0) == B200_OK) { p = dp.t; have = true; }
    }
    if (!have) {
        int rc = fill_params(g, &p, 0);
        if (rc != B200_OK) return rc;
    }
    int n = 0;
    for (int r0 = g->p0; r0 < g->p1; r0 += p.PR) {
        const int r1 = r0 + p.PR < g->p1 ? r0 + p.PR : g->p1;
        if (hi <= lo || (r0 < hi && r1 > lo)) ++n;
    }
    return n * p.tiles_c;
}

extern "C" int b200_tc_block_tiles(const b200_block_geom* g) {
    if (!g) return B200_ERR_ARG;
    if (g->p1 <= g->p0) return 0;
    if (db_requested(g) || pair_requested(g)) {        // same tiling the launch will use
        DbParams dp;
        if (fill_db(g, &dp, 0) == B200_OK) return dp.t.tiles_r * dp.t.tiles_c;
    }
    TcParams p;
    int rc = fill_params(g, &p, 0);
    if (rc != B200_OK) return rc;
    return p.tiles_r * p.tiles_c;
}

// flags bit 0: round the stored outputs to TF32 (use for activations feeding another tc block)
// flags bit 1: the kernel may become resident before the previous kernel in `stream` has finished (programmatic
//              dependent launch); it waits for that kernel's completion before reading anything it wrote.  Honoured
//              by the channel-split kernel, ignored by the others (they launch with the normal stream order).
extern "C" int b200_tc_block(void* handle, const float* bias, float* out, const b200_block_geom* g,
                             int flags, void* stream) {
    TcConv* t = (TcConv*)handle;
    if (!t || !bias || !out || !g) return B200_ERR_ARG;
    if (g->nseg < 1 || g->nseg > B200_MAX_SEG || g->n_signal < 0 || g->n_signal > B200_MAX_SIGNAL ||
        g->n_done < 0 || g->n_done > 2 || g->n_mirror < 0 || g->n_mirror > B200_MAX_MIRROR)
        return B200_ERR_ARG;
    if (g->K != t->K || g->C != t->C || g->F != t->F) return B200_ERR_ARG;
    if (g->p1 <= g->p0 || g->N <= 0) return B200_OK;
    const int sms = t->sms;
    if (t->patch_mode && g->K <= 128 && g->lrn.N_lrn == 0 && !policy().no_swap) {
        SwParams sp;
        memset(&sp, 0, sizeof(sp));
        if (fill_swapped(g, &sp, flags) == B200_OK) {       // else fall through to the generic kernel
            const unsigned grid = (unsigned)(sp.t.total_tiles < sms ? sp.t.total_tiles : sms);
            PixMaps pm;
            memset(&pm, 0, sizeof(pm));
            const bool row_order = sp.direct && t->nkb_tma > 0;     // the handle has weights in the row K order
            if (row_order) {
                sp.direct = 2;
                sp.nkb_main = g->F;
                sp.t.nkb = t->nkb_tma;
            }
            const bool pix_tma = row_order && try_pixel_tma(t, g, &sp, &pm);   // may deepen the ring: size smem after it
            const size_t smem = smem_bytes_swapped(sp);
            if (pix_tma) {
                B200_CUDA_TRY(cudaFuncSetAttribute(tc_block_swapped_kernel<true>,
                                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                tc_block_swapped_kernel<true><<<grid, kSwThreads, smem, (cudaStream_t)stream>>>(t->tmap_tma, sp, bias,
                                                                                               out, pm);
            } else {
                B200_CUDA_TRY(cudaFuncSetAttribute(tc_block_swapped_kernel<false>,
                                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                tc_block_swapped_kernel<false><<<grid, kSwThreads, smem, (cudaStream_t)stream>>>(
                    row_order ? t->tmap_tma : t->tmap, sp, bias, out, pm);
            }
            B200_CUDA_TRY(cudaGetLastError());
            return B200_OK;
        }
    }
    if (cs_requested(g)) {
        CsParams cp;
        CUtensorMap tmapX;
        if (fill_cs(t, g, &cp, &tmapX, flags) == B200_OK) {     // else fall through to the pair / db kernels
            const size_t smem = smem_bytes_cs(cp);
            B200_CUDA_TRY(cudaFuncSetAttribute(tc_block_cs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            long long ctas = 2 * cp.t.total_tiles;
            if (ctas > (sms & ~1)) ctas = sms & ~1;
            cudaLaunchConfig_t cfg;
            memset(&cfg, 0, sizeof(cfg));
            cfg.gridDim = dim3((unsigned)ctas);
            cfg.blockDim = dim3(kCsThreads);
            cfg.dynamicSmemBytes = smem;
            cfg.stream = (cudaStream_t)stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = ((flags & 2) && policy().pdl) ? 1 : 0;
            B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, tc_block_cs_kernel, t->tmap_half, tmapX, cp, bias, (float*)out));
            B200_CUDA_TRY(cudaGetLastError());
            return B200_OK;
        }
    }
    if (pair_requested(g) || cs_requested(g)) {
        PairParams pp;
        CUtensorMap tmapA;
        memset(&pp, 0, sizeof(pp));
        if (fill_pair(t, g, &pp, &tmapA, flags) == B200_OK) {   // else fall through
            const size_t smem = smem_bytes_pair(pp);
            B200_CUDA_TRY(cudaFuncSetAttribute(tc_block_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (int)smem));
            long long ctas = 2 * ((pp.t.total_tiles + 1) / 2);
            if (ctas > (sms & ~1)) ctas = sms & ~1;
            tc_block_pair_kernel<<<(unsigned)ctas, kPairThreads, smem, (cudaStream_t)stream>>>(t->tmap_half, tmapA, pp,
                                                                                              bias, out);
            B200_CUDA_TRY(cudaGetLastError());
            return B200_OK;
        }
    }
    if (db_requested(g)) {
        DbParams dp;
        memset(&dp, 0, sizeof(dp));
        if (fill_db(g, &dp, flags) == B200_OK) {            // else fall through to the single-buffered kernel
            const size_t smem = smem_bytes_db(dp);
            B200_CUDA_TRY(cudaFuncSetAttribute(tc_block_db_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (int)smem));
            const unsigned grid = (unsigned)(dp.t.total_tiles < sms ? dp.t.total_tiles : sms);
            tc_block_db_kernel<<<grid, kDbThreads, smem, (cudaStream_t)stream>>>(t->tmap, dp, bias, out);
            B200_CUDA_TRY(cudaGetLastError());
            return B200_OK;
        }
    }
    TcParams p;
    int rc = fill_params(g, &p, flags);
    if (rc != B200_OK) return rc;
    const size_t smem = smem_bytes(p);
    if (smem > 227 * 1024) return B200_ERR_UNSUPPORTED;
    const unsigned grid = (unsigned)(p.total_tiles < sms ? p.total_tiles : sms);
    if (t->patch_mode) {
        B200_CUDA_TRY(cudaFuncSetAttribute(tc_block_kernel<true>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        tc_block_kernel<true><<<grid, kThreads, smem, (cudaStream_t)stream>>>(t->tmap, p, bias, out);
    } else {
        B200_CUDA_TRY(cudaFuncSetAttribute(tc_block_kernel<false>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        tc_block_kernel<false><<<grid, kThreads, smem, (cudaStream_t)stream>>>(t->tmap, p, bias, out);
    }
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}
