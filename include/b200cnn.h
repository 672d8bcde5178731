/* b200cnn.h — C ABI of the native layers of the b200cnn framework.
 *
 * Two shared objects implement this header:
 *   libb200cnn_cpu.so   csrc/cpu/ (g++)            — oracle, V1/V2 CPU paths, HW1 serial check
 *   libb200cnn_cuda.so  csrc/cuda/ (nvcc sm_100a)  — SIMT exact-FP32 and tcgen05 TF32 kernels,
 *                                                     peer-memory (NVLink) scatter/halo/gather
 * Python binds them with ctypes (b200cnn.ops); C/C++ callers include this file.
 *
 * Tensors are NHWC float32 ("HWC" in the reference, plus a leading batch dimension the
 * reference does not have).  Conv weights are [K][C][F][F] exactly as in
 * ref final_project/v1_serial/src/layers_serial.cpp:70; the CUDA side keeps a re-packed copy.
 *
 * Row slabs: a rank holds rows [base, base+slab) of a logical tensor of full height H.
 * Element (n,h,w,c) of a slab lives at ((n*slab + (h-base))*W + w)*C + c.
 * Every function returns 0 on success, a negative b200_status otherwise.
 */
#ifndef B200CNN_H
#define B200CNN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    B200_OK = 0,
    B200_ERR_ARG = -1,       /* bad geometry / null pointer                      */
    B200_ERR_CUDA = -2,      /* CUDA runtime error; see b200_last_error()        */
    B200_ERR_RANGE = -3,     /* a needed input row is outside the supplied slab  */
    B200_ERR_TIMEOUT = -4,   /* a cross-GPU flag wait hit the watchdog           */
    B200_ERR_UNSUPPORTED = -5
} b200_status;

/* Window-op geometry shared by conv and pool (CPU and CUDA). */
typedef struct {
    int N;                   /* images                                                   */
    int H, W, C;             /* logical input height / width / channels                  */
    int in_base, in_slab;    /* input slab: global row of slab row 0, rows per image     */
    int K;                   /* output channels (pool: K == C)                           */
    int F, S, P;             /* window, stride, zero padding (pool: P == 0)              */
    int Ho, Wo;              /* logical output height / width                            */
    int out_base, out_slab;  /* output slab                                              */
    int o0, o1;              /* global output rows to compute, half-open                 */
} b200_win_geom;

/* LRN conventions: the reference disagrees with itself. */
enum { B200_LRN_ALPHA_OVER_N = 0 /* V1/V2 */, B200_LRN_ALPHA_RAW = 1 /* V3/V4 */ };

typedef struct {
    int N_lrn;               /* window (0 => copy)            */
    float alpha, beta, k;
    int mode;                /* B200_LRN_*                    */
} b200_lrn_params;

/* ------------------------------------------------------------------ CPU library ---- */

/* ref v1_serial/src/layers_serial.cpp:37-81 (direct conv, bias first, zero pad by bounds check),
 * summation order per output = (c, fh, fw) like the reference, so results are bit-comparable. */
int b200_cpu_conv(const float* in, const float* w, const float* bias, float* out,
                  const b200_win_geom* g);
/* ref layers_serial.cpp:85-90 */
int b200_cpu_relu(float* data, size_t n);
/* ref layers_serial.cpp:94-129 (init -inf, window clipped to the tensor) */
int b200_cpu_maxpool(const float* in, float* out, const b200_win_geom* g);
/* ref layers_serial.cpp:133-175; npix = N*H*W pixels of C channels */
int b200_cpu_lrn(const float* in, float* out, size_t npix, int C, const b200_lrn_params* p);

/* glibc rand() TYPE_3 generator reproduced so a seed gives the V1 stream on any libc
 * (ref v1_serial/src/alexnet_serial.cpp:39-57; draw order main.cpp:51-62). state: 34+1 words. */
typedef struct { int32_t r[34]; int32_t idx; } b200_grand;
void b200_grand_seed(b200_grand* st, uint32_t seed);
int32_t b200_grand_next(b200_grand* st);
void b200_grand_fill_data(b200_grand* st, float* out, size_t n);      /* rand/RAND_MAX*0.1       */
void b200_grand_fill_weights(b200_grand* st, float* out, size_t n);   /* (rand/RAND_MAX-.5)*.02  */

/* FNV-1a 64 over raw bytes: the output digest summary.md:108 describes but never implements. */
uint64_t b200_fnv1a64(const void* data, size_t nbytes);

/* HW1 (ref homeworks/hw1/src/template.c:195-238): C = A*B row block, init, check. */
void b200_cpu_matmul_f64(const double* A, const double* B, double* C, int rows, int n);
void b200_hw1_init_data(double* M, int rows, int cols);   /* fixed srand48(1) stream like init_data */
int  b200_hw1_check(const double* C, const double* Cref, size_t n, double tol);

/* ----------------------------------------------------------------- CUDA library ---- */

const char* b200_last_error(void);
int b200_cuda_device_count(void);

/* Exact-FP32 SIMT path (V3/V4 kernel equivalents, fused; ref v3_cuda_only/src/layers_cuda.cu:20-139,
 * v4_mpi_cuda/src/layers_mpi_cuda.cu:25-89).  conv + bias + ReLU + max-pool in one kernel;
 * optionally + LRN.  `g` describes the conv; pool geometry is (Fp,Sp) on the conv output;
 * pooled rows [p0,p1) are written to `out` slab (pool_base, pool_slab).
 * wt = weights re-packed [F][F][C][K] (b200_pack_weights_fck). */
typedef struct {
    b200_win_geom conv;      /* o0/o1 ignored: derived from pooled rows                   */
    int Fp, Sp;              /* pool window / stride                                      */
    int Hp, Wp;              /* logical pooled dims                                       */
    int pool_base, pool_slab;
    int p0, p1;              /* pooled rows to produce                                    */
    b200_lrn_params lrn;     /* N_lrn == 0 => no LRN                                      */
} b200_block_geom;

int b200_pack_weights_kfc(const float* w_kcff, float* w_kffc, int K, int C, int F);   /* host */

int b200_simt_block(const float* in, const float* w_kffc, const float* bias, float* out,
                    const b200_block_geom* g, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200CNN_H */
