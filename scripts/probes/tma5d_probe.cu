// Probe: does a 5-D TMA box {4, 8, TC, TR, 1} over an overlapping-stride view of an NHWC-pitched image land as
// 128 B pixel rows (chunk j = input row 8t + j, floats 32..35 of the run) in the 128B-swizzled layout?
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tma5d_probe tma5d_probe.cu -lcuda ; ./tma5d_probe [promo] [t]
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cstdint>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); return 2; } } while (0)

__global__ void probe(const __grid_constant__ CUtensorMap map, float* out, int x0, int y0, int n, int bytes) {
    extern __shared__ uint8_t raw[];
    const uint32_t base = ((uint32_t)__cvta_generic_to_shared(raw) + 1023u) & ~1023u;
    uint8_t* gb = raw + (base - (uint32_t)__cvta_generic_to_shared(raw));
    const uint32_t bar = base + 40960;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < 40960 / 4; i += blockDim.x) ((float*)gb)[i] = -7.f;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile(
            "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
            ::"r"(base), "l"(&map), "r"(0), "r"(0), "r"(x0), "r"(y0), "r"(n), "r"(bar) : "memory");
    }
    uint32_t done = 0;
    for (long it = 0; !done && it < 4000000; ++it)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(bar), "r"(0) : "memory");
    __syncthreads();
    for (int i = threadIdx.x; i < 40960 / 4; i += blockDim.x) out[i] = ((float*)gb)[i];
    if (threadIdx.x == 0) out[40960 / 4] = done ? 1.f : 0.f;
}

int main(int argc, char** argv) {
    const int promo = argc > 1 ? atoi(argv[1]) : 2, t = argc > 2 ? atoi(argv[2]) : 0;
    const int N = 2, H = 227, W = 227, C = 3, pitch = 684, S = 4, F = 11, Wc = 55, TC = 13, TR = 19;
    std::vector<float> h((size_t)N * H * pitch);
    for (int n = 0; n < N; ++n) for (int r = 0; r < H; ++r) for (int c = 0; c < pitch; ++c)
        h[((size_t)n * H + r) * pitch + c] = (float)(n * 1000000 + r * 1000 + c);
    float *d, *o;
    CK(cudaMalloc(&d, h.size() * 4)); CK(cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&o, 40960 + 64));
    typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                            const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                            CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    const int J = F - 8 * t < 8 ? F - 8 * t : 8, slab = H;
    const int rows = (slab - 8 * t - J) / S + 1;
    cuuint64_t dims[5] = {4, (cuuint64_t)J, (cuuint64_t)Wc, (cuuint64_t)rows, (cuuint64_t)N};
    cuuint64_t strides[4] = {(cuuint64_t)pitch * 4, (cuuint64_t)S * C * 4, (cuuint64_t)S * pitch * 4, (cuuint64_t)slab * pitch * 4};
    cuuint32_t box[5] = {4, 8, (cuuint32_t)TC, (cuuint32_t)TR, 1}, estr[5] = {1, 1, 1, 1, 1};
    CUtensorMap map;
    CUresult rc = ((Enc)fn)(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 5, d + (size_t)8 * t * pitch + 32, dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, (CUtensorMapL2promotion)promo,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d (promo %d, tail %d, J %d, rows %d)\n", (int)rc, promo, t, J, rows);
    if (rc != CUDA_SUCCESS) return 1;
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024));
    int bad_total = 0;
    const int cases[3][3] = {{0, 0, 0}, {26, 20, 1}, {52, 38, 1}};      // (x0, y0, n): interior and the bottom-right corner
    for (int cs = 0; cs < 3; ++cs) {
        const int x0 = cases[cs][0], y0 = cases[cs][1], n = cases[cs][2];
        probe<<<1, 256, 48 * 1024>>>(map, o, x0, y0, n, 4 * 8 * TC * TR * 4);
        CK(cudaDeviceSynchronize());
        std::vector<float> r(40960 / 4 + 1);
        CK(cudaMemcpy(r.data(), o, 40960 + 4, cudaMemcpyDeviceToHost));
        int bad = 0;
        for (int pi = 0; pi < TR; ++pi) for (int pj = 0; pj < TC; ++pj) for (int j = 0; j < 8; ++j) for (int i = 0; i < 4; ++i) {
            const int m = pi * TC + pj;
            const float got = r[(size_t)m * 32 + ((j ^ (m & 7)) << 2) + i];
            const int x = x0 + pj, y = y0 + pi, row = S * y + 8 * t + j;
            const bool in = j < J && x < Wc && y < rows;
            const float want = in ? (float)(n * 1000000 + row * 1000 + x * S * C + 32 + i) : 0.f;
            if (got != want && bad++ < 5) printf("  case %d slot %d chunk %d float %d: got %.0f want %.0f\n", cs, m, j, i, got, want);
        }
        printf("case %d (x0 %d, y0 %d, n %d): barrier %s, %d mismatches\n", cs, x0, y0, n, r[40960 / 4] == 1.f ? "completed" : "TIMED OUT", bad);
        bad_total += bad;
    }
    printf(bad_total ? "PROBE FAILED\n" : "PROBE OK\n");
    return bad_total ? 1 : 0;
}
