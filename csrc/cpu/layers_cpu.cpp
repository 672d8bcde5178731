// CPU layer ops: the numeric oracle and the compute of the V1 / V2.1 / V2.2 paths.
//
// Math follows ref final_project/v1_serial/src/layers_serial.cpp:37-175 (conv: bias first,
// taps summed in (c, fh, fw) order, zero padding by bounds check; pool: init -inf, window
// clipped; LRN: clamped channel window).  Unlike the reference these ops (a) carry a batch
// dimension, (b) work on row slabs so the same code serves the exact spatial decomposition,
// (c) keep the reference's per-output summation order while vectorising ACROSS output
// channels (weights transposed to [tap][K]), so results stay bit-comparable with V1 built
// without FMA contraction.
#include "b200cnn.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

namespace {

inline bool geom_ok(const b200_win_geom* g) {
    return g && g->N >= 0 && g->H > 0 && g->W > 0 && g->C > 0 && g->K > 0 && g->F > 0 &&
           g->S > 0 && g->P >= 0 && g->Ho >= 0 && g->Wo >= 0 && g->o0 <= g->o1 &&
           g->in_slab >= 0 && g->out_slab >= 0;
}

// acc[k] += x * wt[k] for k in [0,K): the only hot loop.  target_clones lets the loader pick
// AVX2 when the host has it; -ffp-contract=off (build flag) keeps mul and add separate.
__attribute__((target_clones("avx2", "default")))
void axpy_k(float* __restrict__ acc, const float* __restrict__ wt, float x, int K) {
    for (int k = 0; k < K; ++k) acc[k] += x * wt[k];
}

}  // namespace

extern "C" int b200_cpu_conv(const float* in, const float* w, const float* bias, float* out,
                             const b200_win_geom* g) {
    if (!geom_ok(g) || !in || !w || !bias || !out) return B200_ERR_ARG;
    const int C = g->C, K = g->K, F = g->F, S = g->S, P = g->P;
    const int taps = C * F * F;
    // [K][C][F][F] -> [C][F][F][K]; tap order (c, fh, fw) is the reference's summation order.
    std::vector<float> wt((size_t)taps * K);
    for (int k = 0; k < K; ++k)
        for (int t = 0; t < taps; ++t) wt[(size_t)t * K + k] = w[(size_t)k * taps + t];

    // every needed in-range input row must be inside the slab
    if (g->o1 > g->o0) {
        int lo = std::max(g->o0 * S - P, 0), hi = std::min((g->o1 - 1) * S - P + F, g->H);
        if (lo < g->in_base || hi > g->in_base + g->in_slab) return B200_ERR_RANGE;
        if (g->o0 < g->out_base || g->o1 > g->out_base + g->out_slab) return B200_ERR_RANGE;
    }
    const size_t in_img = (size_t)g->in_slab * g->W * C;
    const size_t out_img = (size_t)g->out_slab * g->Wo * K;
    const int rows = g->o1 - g->o0;
#pragma omp parallel for collapse(2) schedule(static)
    for (int n = 0; n < g->N; ++n) {
        for (int r = 0; r < rows; ++r) {
            const int ho = g->o0 + r;
            const float* img = in + (size_t)n * in_img;
            float* orow = out + (size_t)n * out_img + (size_t)(ho - g->out_base) * g->Wo * K;
            for (int wo = 0; wo < g->Wo; ++wo) {
                float* acc = orow + (size_t)wo * K;
                for (int k = 0; k < K; ++k) acc[k] = bias[k];
                for (int c = 0; c < C; ++c)
                    for (int fh = 0; fh < F; ++fh) {
                        const int hi = ho * S - P + fh;
                        if (hi < 0 || hi >= g->H) continue;
                        const float* irow = img + (size_t)(hi - g->in_base) * g->W * C;
                        for (int fw = 0; fw < F; ++fw) {
                            const int wi = wo * S - P + fw;
                            if (wi < 0 || wi >= g->W) continue;
                            const int t = (c * F + fh) * F + fw;
                            axpy_k(acc, &wt[(size_t)t * K], irow[(size_t)wi * C + c], K);
                        }
                    }
            }
        }
    }
    return B200_OK;
}

extern "C" int b200_cpu_relu(float* data, size_t n) {
    if (!data && n) return B200_ERR_ARG;
    for (size_t i = 0; i < n; ++i) data[i] = std::max(0.0f, data[i]);
    return B200_OK;
}

extern "C" int b200_cpu_maxpool(const float* in, float* out, const b200_win_geom* g) {
    if (!geom_ok(g) || !in || !out || g->K != g->C) return B200_ERR_ARG;
    const int C = g->C, F = g->F, S = g->S;
    if (g->o1 > g->o0) {
        int lo = g->o0 * S, hi = std::min((g->o1 - 1) * S + F, g->H);
        if (lo < g->in_base || hi > g->in_base + g->in_slab) return B200_ERR_RANGE;
        if (g->o0 < g->out_base || g->o1 > g->out_base + g->out_slab) return B200_ERR_RANGE;
    }
    const size_t in_img = (size_t)g->in_slab * g->W * C;
    const size_t out_img = (size_t)g->out_slab * g->Wo * C;
#pragma omp parallel for schedule(static)
    for (int n = 0; n < g->N; ++n)
        for (int ho = g->o0; ho < g->o1; ++ho)
            for (int wo = 0; wo < g->Wo; ++wo) {
                float* o = out + (size_t)n * out_img + ((size_t)(ho - g->out_base) * g->Wo + wo) * C;
                for (int c = 0; c < C; ++c) o[c] = -std::numeric_limits<float>::infinity();
                for (int fh = 0; fh < F; ++fh) {
                    const int hi = ho * S + fh;
                    if (hi >= g->H) continue;
                    for (int fw = 0; fw < F; ++fw) {
                        const int wi = wo * S + fw;
                        if (wi >= g->W) continue;
                        const float* p = in + (size_t)n * in_img +
                                         ((size_t)(hi - g->in_base) * g->W + wi) * C;
                        for (int c = 0; c < C; ++c) o[c] = std::max(o[c], p[c]);
                    }
                }
            }
    return B200_OK;
}

extern "C" int b200_cpu_lrn(const float* in, float* out, size_t npix, int C,
                            const b200_lrn_params* p) {
    if (!in || !out || !p || C <= 0) return B200_ERR_ARG;
    if (p->N_lrn <= 0) {  // ref layers_serial.cpp:146-151: copy through
        std::memcpy(out, in, npix * C * sizeof(float));
        return B200_OK;
    }
    const int half = p->N_lrn / 2;
    const float a = p->mode == B200_LRN_ALPHA_OVER_N ? p->alpha / (float)p->N_lrn : p->alpha;
#pragma omp parallel for schedule(static)
    for (long long px = 0; px < (long long)npix; ++px) {
        const float* x = in + (size_t)px * C;
        float* y = out + (size_t)px * C;
        for (int c = 0; c < C; ++c) {
            float ss = 0.0f;
            const int c0 = std::max(0, c - half), c1 = std::min(C - 1, c + half);
            for (int i = c0; i <= c1; ++i) ss += x[i] * x[i];
            y[c] = x[c] / std::pow(p->k + a * ss, p->beta);
        }
    }
    return B200_OK;
}
