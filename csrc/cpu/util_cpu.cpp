// Host utilities: glibc-compatible rand() stream, FNV-1a digest, HW1 serial helpers.
#include "b200cnn.h"

#include <cmath>
#include <cstdio>
#include <cstring>

// ---- glibc random_r TYPE_3 (degree 31, separation 3) ---------------------------------------
// srand(seed): r[0]=seed (0 -> 1); r[i] = 16807*r[i-1] mod (2^31-1) for i<31 (Schrage, sign
// fixed); 310 outputs discarded.  rand(): r[k] = r[k-31] + r[k-3] (mod 2^32), result >> 1.
// Keeping our own copy makes `--init rand --seed S` give V1's stream on any libc and from
// any thread (ref v1_serial/src/main.cpp:12, alexnet_serial.cpp:39-57).
extern "C" void b200_grand_seed(b200_grand* st, uint32_t seed) {
    int32_t r[34];
    if (seed == 0) seed = 1;
    r[0] = (int32_t)seed;
    for (int i = 1; i < 31; ++i) {
        long hi = r[i - 1] / 127773, lo = r[i - 1] % 127773;
        long word = 16807 * lo - 2836 * hi;
        if (word < 0) word += 2147483647;
        r[i] = (int32_t)word;
    }
    // ring buffer of the last 31 values; warm up by discarding 310 draws
    for (int i = 0; i < 31; ++i) st->r[i] = r[i];
    st->r[31] = st->r[32] = st->r[33] = 0;
    st->idx = 0;
    // f = idx+3 (mod 31), r = idx: glibc keeps fptr = &state[3], rptr = &state[0]
    for (int i = 0; i < 310; ++i) (void)b200_grand_next(st);
}

extern "C" int32_t b200_grand_next(b200_grand* st) {
    const int rp = st->idx, fp = (st->idx + 3) % 31;
    uint32_t v = (uint32_t)st->r[fp] + (uint32_t)st->r[rp];
    st->r[fp] = (int32_t)v;
    st->idx = (st->idx + 1) % 31;
    return (int32_t)(v >> 1);
}

extern "C" void b200_grand_fill_data(b200_grand* st, float* out, size_t n) {
    const float rmax = (float)2147483647;
    for (size_t i = 0; i < n; ++i) out[i] = (float)b200_grand_next(st) / rmax * 0.1f;
}

extern "C" void b200_grand_fill_weights(b200_grand* st, float* out, size_t n) {
    const float rmax = (float)2147483647;
    for (size_t i = 0; i < n; ++i) out[i] = ((float)b200_grand_next(st) / rmax - 0.5f) * 0.02f;
}

extern "C" uint64_t b200_fnv1a64(const void* data, size_t nbytes) {
    const unsigned char* p = (const unsigned char*)data;
    uint64_t h = 14695981039346656037ULL;
    for (size_t i = 0; i < nbytes; ++i) { h ^= p[i]; h *= 1099511628211ULL; }
    return h;
}

// ---- HW1 (ref homeworks/hw1/src/template.c:195-238) ----------------------------------------
extern "C" void b200_cpu_matmul_f64(const double* A, const double* B, double* C, int rows, int n) {
    // i-k-j order: same products, each C[i][j] still accumulates k ascending (bit-equal to the
    // reference's i-j-k dot product) but streams B rows, which is what makes n=2048 tolerable.
#pragma omp parallel for schedule(static)
    for (int i = 0; i < rows; ++i) {
        double* c = C + (size_t)i * n;
        for (int j = 0; j < n; ++j) c[j] = 0.0;
        for (int k = 0; k < n; ++k) {
            const double a = A[(size_t)i * n + k];
            const double* b = B + (size_t)k * n;
            for (int j = 0; j < n; ++j) c[j] += a * b[j];
        }
    }
}

extern "C" void b200_hw1_init_data(double* M, int rows, int cols) {
    // integers 0..9 as doubles (template.c:211-216); the stream is ours (xorshift) because the
    // reference seeds with time(NULL) and so fixes no particular values.
    static uint64_t s = 0x9E3779B97F4A7C15ULL;
    for (size_t i = 0; i < (size_t)rows * cols; ++i) {
        s ^= s << 13; s ^= s >> 7; s ^= s << 17;
        M[i] = (double)(s % 10);
    }
}

extern "C" int b200_hw1_check(const double* C, const double* Cref, size_t n, double tol) {
    int flag = 0;
    for (size_t i = 0; i < n; ++i) {
        double d = C[i] - Cref[i];
        if (d > tol || d < -tol || d != d) flag = 1;
    }
    return flag;
}
