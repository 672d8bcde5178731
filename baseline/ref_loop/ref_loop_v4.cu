// Warm loop around the reference's unmodified V4 tile forward (alexnetTileForwardCUDA,
// ref final_project/v4_mpi_cuda/src/alexnet_mpi_cuda.cu:157-205) on one GPU with the whole 227x227x3 image as the
// tile (what V4 runs at np 1): constant initialisation as in ref v4_mpi_cuda/src/main_mpi_cuda.cpp:27-34, the image
// resident on the device, W warm-up calls, then N timed calls = N images.  Every call re-uploads the 2.6 MB of
// weights and allocates/frees its 8 buffers, as the reference wrote it; start-up, context creation and the MPI
// scatter/gather are outside the loop.
//   ref_loop_v4 [images=128] [warmup=3]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include <mpi.h>
#include "alexnet.hpp"

int main(int argc, char** argv) {
    MPI_Init(&argc, &argv);
    const int images = argc > 1 ? atoi(argv[1]) : 128, warm = argc > 2 ? atoi(argv[2]) : 3;
    const int H = 227, W = 227, C = 3;
    LayerParams p1, p2;
    p1.K = 96; p1.F = 11; p1.S = 4; p1.P = 0; p1.F_pool = 3; p1.S_pool = 2; p1.N_lrn = 0;
    p1.alpha = 0; p1.beta = 0; p1.k_lrn = 0;
    p1.weights.assign((size_t)p1.K * C * p1.F * p1.F, 0.01f); p1.biases.assign(p1.K, 0.0f);
    p2.K = 256; p2.F = 5; p2.S = 1; p2.P = 2; p2.F_pool = 3; p2.S_pool = 2; p2.N_lrn = 5;
    p2.alpha = 1e-4f; p2.beta = 0.75f; p2.k_lrn = 2.0f;
    p2.weights.assign((size_t)p2.K * p1.K * p2.F * p2.F, 0.01f); p2.biases.assign(p2.K, 0.0f);
    std::vector<float> in((size_t)H * W * C, 1.0f), out((size_t)13 * 13 * 256);
    float *d_in = nullptr, *d_out = nullptr;
    cudaMalloc(&d_in, in.size() * sizeof(float));
    cudaMalloc(&d_out, out.size() * sizeof(float));
    cudaMemcpy(d_in, in.data(), in.size() * sizeof(float), cudaMemcpyHostToDevice);
    for (int i = 0; i < warm; ++i) alexnetTileForwardCUDA(d_in, p1, p2, H, W, C, d_out);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto t0 = std::chrono::steady_clock::now();
    cudaEventRecord(e0);
    for (int i = 0; i < images; ++i) alexnetTileForwardCUDA(d_in, p1, p2, H, W, C, d_out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    auto t1 = std::chrono::steady_clock::now();
    float dev_ms = 0;
    cudaEventElapsedTime(&dev_ms, e0, e1);
    const double wall_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    cudaMemcpy(out.data(), d_out, out.size() * sizeof(float), cudaMemcpyDeviceToHost);
    printf("ref_loop_v4: images=%d warmup=%d device_ms_total=%.3f wall_ms_total=%.3f ms_per_image=%.4f images_per_s=%.1f "
           "out[0]=%.4f\n", images, warm, dev_ms, wall_ms, dev_ms / images, images / (dev_ms / 1e3), out[0]);
    cudaFree(d_in); cudaFree(d_out);
    MPI_Finalize();
    return 0;
}
