// Warm loop around the reference's unmodified V3 forward (alexnetForwardPassCUDA,
// ref final_project/v3_cuda_only/src/alexnet_cuda.cu:22-95): same constant initialisation as the reference's
// main (ref v3_cuda_only/src/main_cuda.cpp:14-27), W warm-up calls, then N timed calls = N images (the reference
// has no batch dimension).  The function itself allocates, uploads, launches its 7 kernels, downloads and frees
// on every call; what the warm loop removes is process start-up and CUDA context creation.
//   ref_loop_v3 [images=128] [warmup=3]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "alexnet.hpp"

int main(int argc, char** argv) {
    const int images = argc > 1 ? atoi(argv[1]) : 128, warm = argc > 2 ? atoi(argv[2]) : 3;
    const int H = 227, W = 227, C = 3;
    std::vector<float> in((size_t)H * W * C, 1.0f), out;
    LayerParams p1, p2;
    p1.K = 96; p1.F = 11; p1.S = 4; p1.P = 0; p1.F_pool = 3; p1.S_pool = 2; p1.N_lrn = 0;
    p1.alpha = 0; p1.beta = 0; p1.k_lrn = 0;
    p1.weights.assign((size_t)p1.K * C * p1.F * p1.F, 0.01f); p1.biases.assign(p1.K, 0.0f);
    p2.K = 256; p2.F = 5; p2.S = 1; p2.P = 2; p2.F_pool = 3; p2.S_pool = 2; p2.N_lrn = 5;
    p2.alpha = 1e-4f; p2.beta = 0.75f; p2.k_lrn = 2.0f;
    p2.weights.assign((size_t)p2.K * p1.K * p2.F * p2.F, 0.01f); p2.biases.assign(p2.K, 0.0f);
    for (int i = 0; i < warm; ++i) alexnetForwardPassCUDA(in, p1, p2, H, W, C, out);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto t0 = std::chrono::steady_clock::now();
    cudaEventRecord(e0);
    for (int i = 0; i < images; ++i) alexnetForwardPassCUDA(in, p1, p2, H, W, C, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    auto t1 = std::chrono::steady_clock::now();
    float dev_ms = 0;
    cudaEventElapsedTime(&dev_ms, e0, e1);
    const double wall_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    printf("ref_loop_v3: images=%d warmup=%d device_ms_total=%.3f wall_ms_total=%.3f ms_per_image=%.4f images_per_s=%.1f "
           "out[0]=%.4f out_size=%zu\n", images, warm, dev_ms, wall_ms, dev_ms / images, images / (dev_ms / 1e3),
           out.empty() ? 0.f : out[0], out.size());
    return 0;
}
