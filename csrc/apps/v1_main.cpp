// Native V1 `template`: serial CPU forward of AlexNet blocks 1-2, printing the reference's exact
// stdout contract (ref final_project/v1_serial/src/main.cpp:10-78, alexnet_serial.cpp:61-185).
// Built by final_project/v1_serial/Makefile against libb200cnn_cpu.so (include/b200cnn.h).
//
// Zero arguments == reference behaviour (rand() stream seeded with time(0), biases 0.1, alpha/N
// LRN).  Additions: --seed S pins the stream; --quiet-extras prints only the reference's lines.
#include "b200cnn.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <iostream>
#include <vector>

static void dims(const char* layer, int H, int W, int C) {
    std::cout << "  [" << layer << "] Dimensions: H=" << H << ", W=" << W << ", C=" << C << std::endl;
}

int main(int argc, char** argv) {
    unsigned seed = (unsigned)time(nullptr);
    bool extras = true;
    const char* load_weights = nullptr;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--seed") && i + 1 < argc) seed = (unsigned)strtoul(argv[++i], nullptr, 10);
        else if (!strcmp(argv[i], "--quiet-extras")) extras = false;
        else if (!strcmp(argv[i], "--load-weights") && i + 1 < argc) load_weights = argv[++i];
    }
    std::cout << "--- AlexNet Serial CPU Version (V1) ---" << std::endl;
    const int H = 227, W = 227, C = 3;
    const int K1 = 96, F1 = 11, S1 = 4, P1 = 0, K2 = 256, F2 = 5, S2 = 1, P2 = 2, Fp = 3, Sp = 2;
    std::cout << "Initializing data and parameters..." << std::endl;
    b200_grand st;
    b200_grand_seed(&st, seed);
    std::vector<float> x((size_t)H * W * C), w1((size_t)K1 * C * F1 * F1), b1(K1, 0.1f),
        w2((size_t)K2 * K1 * F2 * F2), b2(K2, 0.1f);
    b200_grand_fill_data(&st, x.data(), x.size());          // draw order: input, W1, W2
    b200_grand_fill_weights(&st, w1.data(), w1.size());
    b200_grand_fill_weights(&st, w2.data(), w2.size());
    if (load_weights) {                        // parameters from a weight file (utils/checkpoint.py); input as drawn
        struct { const char* name; std::vector<float>* t; } want[] = {{"w1", &w1}, {"b1", &b1}, {"w2", &w2}, {"b2", &b2}};
        for (auto& q : want) {
            size_t n = 0;
            if (b200_weights_read(load_weights, q.name, nullptr, 0, &n, nullptr, nullptr) || n != q.t->size() ||
                b200_weights_read(load_weights, q.name, q.t->data(), q.t->size(), &n, nullptr, nullptr)) {
                std::cerr << "[Rank 0] --load-weights " << load_weights << ": cannot read " << q.name
                          << " (missing file, wrong shape or checksum mismatch)" << std::endl;
                return 1;
            }
        }
    }
    std::cout << "Initialization complete." << std::endl;

    std::cout << "Starting AlexNet Serial Forward Pass..." << std::endl;
    auto t0 = std::chrono::high_resolution_clock::now();
    const int Hc1 = (H - F1 + 2 * P1) / S1 + 1, Wc1 = (W - F1 + 2 * P1) / S1 + 1;
    const int Hp1 = (Hc1 - Fp) / Sp + 1, Wp1 = (Wc1 - Fp) / Sp + 1;
    const int Hc2 = (Hp1 - F2 + 2 * P2) / S2 + 1, Wc2 = (Wp1 - F2 + 2 * P2) / S2 + 1;
    const int Hp2 = (Hc2 - Fp) / Sp + 1, Wp2 = (Wc2 - Fp) / Sp + 1;
    std::vector<float> c1((size_t)Hc1 * Wc1 * K1), p1((size_t)Hp1 * Wp1 * K1),
        c2((size_t)Hc2 * Wc2 * K2), p2((size_t)Hp2 * Wp2 * K2), out(p2.size());
    dims("Input", H, W, C);
    std::cout << "Applying Conv1..." << std::endl;
    b200_win_geom g = {1, H, W, C, 0, H, K1, F1, S1, P1, Hc1, Wc1, 0, Hc1, 0, Hc1};
    if (b200_cpu_conv(x.data(), w1.data(), b1.data(), c1.data(), &g)) return 1;
    dims("After Conv1", Hc1, Wc1, K1);
    std::cout << "Applying ReLU1..." << std::endl;
    b200_cpu_relu(c1.data(), c1.size());
    dims("After ReLU1", Hc1, Wc1, K1);
    std::cout << "Applying MaxPool1..." << std::endl;
    g = {1, Hc1, Wc1, K1, 0, Hc1, K1, Fp, Sp, 0, Hp1, Wp1, 0, Hp1, 0, Hp1};
    if (b200_cpu_maxpool(c1.data(), p1.data(), &g)) return 1;
    dims("After Pool1", Hp1, Wp1, K1);
    std::cout << "Applying Conv2..." << std::endl;
    g = {1, Hp1, Wp1, K1, 0, Hp1, K2, F2, S2, P2, Hc2, Wc2, 0, Hc2, 0, Hc2};
    if (b200_cpu_conv(p1.data(), w2.data(), b2.data(), c2.data(), &g)) return 1;
    dims("After Conv2", Hc2, Wc2, K2);
    std::cout << "Applying ReLU2..." << std::endl;
    b200_cpu_relu(c2.data(), c2.size());
    dims("After ReLU2", Hc2, Wc2, K2);
    std::cout << "Applying MaxPool2..." << std::endl;
    g = {1, Hc2, Wc2, K2, 0, Hc2, K2, Fp, Sp, 0, Hp2, Wp2, 0, Hp2, 0, Hp2};
    if (b200_cpu_maxpool(c2.data(), p2.data(), &g)) return 1;
    dims("After Pool2", Hp2, Wp2, K2);
    std::cout << "Applying LRN2..." << std::endl;
    b200_lrn_params lp = {5, 0.0001f, 0.75f, 2.0f, B200_LRN_ALPHA_OVER_N};
    if (b200_cpu_lrn(p2.data(), out.data(), (size_t)Hp2 * Wp2, K2, &lp)) return 1;
    dims("After LRN2", Hp2, Wp2, K2);
    std::cout << "Copying final result back to original buffer..." << std::endl;
    x = out;
    auto t1 = std::chrono::high_resolution_clock::now();
    auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(t1 - t0).count();
    std::cout << "AlexNet Serial Forward Pass completed in " << ms << " ms" << std::endl;
    std::cout << "Final Output (first 10 values): ";
    for (int i = 0; i < 10; ++i) std::cout << x[i] << (i == 9 ? "" : " ");
    std::cout << "..." << std::endl;
    std::cout << "--- Serial execution finished ---" << std::endl;
    if (extras) {
        printf("Final Output Shape: %dx%dx%d\n", Hp2, Wp2, K2);
        printf("Output FNV1a64: 0x%016llx\n", (unsigned long long)b200_fnv1a64(x.data(), x.size() * 4));
        printf("seed=%u\n", seed);
        printf("ExecutionTime_ms: %g\n",
               std::chrono::duration<double, std::milli>(t1 - t0).count());
    }
    return 0;
}
