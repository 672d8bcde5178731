// Library-level host utilities of libb200cnn_cuda.so: error latch, device query, weight packing.
#include "common.cuh"

#include <cstdarg>
#include <cstring>
#include <mutex>

namespace b200 {
static std::mutex g_err_mu;
static char g_err[1024] = "";

void set_error(const char* fmt, ...) {
    std::lock_guard<std::mutex> lk(g_err_mu);
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace b200

extern "C" const char* b200_last_error(void) { return b200::g_err; }

extern "C" int b200_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

// [K][C][F][F] (ref v1_serial/src/layers_serial.cpp:70) -> [F][F][C][K]: the (fx,c) run of one
// filter row becomes contiguous, matching the NHWC input run it multiplies.
extern "C" int b200_pack_weights_ffck(const float* w, float* o, int K, int C, int F) {
    if (!w || !o || K <= 0 || C <= 0 || F <= 0) return B200_ERR_ARG;
    for (int k = 0; k < K; ++k)
        for (int c = 0; c < C; ++c)
            for (int fy = 0; fy < F; ++fy)
                for (int fx = 0; fx < F; ++fx)
                    o[(((size_t)fy * F + fx) * C + c) * K + k] =
                        w[(((size_t)k * C + c) * F + fy) * F + fx];
    return B200_OK;
}
