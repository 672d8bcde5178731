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

/* Exact row (x batch) decomposition of conv -> pool -> conv -> pool (csrc/cpu/planner_cpu.cpp; the Python planner
 * pkg/parallel/planner.py is its twin).  Final rows are split base+remainder over the `row_groups` ranks of a group
 * (0 = all ranks: the reference's pure row split), every group takes its share of the batch, and each earlier layer's
 * row range follows from the receptive field.  strategy 0 = halo (pool1 rows disjointly owned, the rest read from / pushed
 * by the peers listed in halo[] / serves[]), 1 = recompute (every rank computes all pool1 rows it needs).  Ranges are
 * half-open; an empty range is (x, x). */
typedef struct { int F, S, P, F_pool, S_pool; } b200_plan_block;
typedef struct { int peer, lo, hi; } b200_plan_halo;
#define B200_PLAN_MAX_HALO 8
typedef struct {
    int rank, world, row_group, batch_group, row_groups, batch_groups;
    int img0, img1;          /* images of the global batch this rank works on               */
    int out0, out1;          /* owned final rows (pool2 / LRN2)                             */
    int conv2_0, conv2_1;    /* conv2 rows computed                                         */
    int need0, need1;        /* pool1 rows conv2 reads (clipped; outside = zero padding)    */
    int own0, own1;          /* pool1 rows produced                                         */
    int conv1_0, conv1_1;    /* conv1 rows computed for them                                */
    int in0, in1;            /* input rows conv1 reads: the rank's "scatter" slab           */
    int n_halo, n_serves;
    b200_plan_halo halo[B200_PLAN_MAX_HALO];     /* pool1 rows [lo,hi) read from `peer`     */
    b200_plan_halo serves[B200_PLAN_MAX_HALO];   /* pool1 rows [lo,hi) `peer` reads from me */
    int Hc1, Hp1, Hc2, Hp2;  /* full heights of the four layers                             */
} b200_rank_plan;
int b200_plan_rank(int H, int batch, const b200_plan_block* b1, const b200_plan_block* b2, int world,
                   int strategy, int row_groups, int rank, b200_rank_plan* out);

/* Weight files written by pkg/utils/checkpoint.py (save_weights): read tensor `name` ("w1", "b1", "w2", "b2", and
 * "w3".."b8" in full-network files) into dst[capacity].  *count = its float count (also with dst == NULL, to size a
 * buffer); shape[0..*ndim) when both are non-NULL (at most 4 dims).  B200_ERR_ARG: file cannot be opened;
 * B200_ERR_UNSUPPORTED: not a weight file or no such tensor; B200_ERR_RANGE: buffer too small, truncated file, or the
 * tensor's FNV-1a64 does not match the header. */
int b200_weights_read(const char* path, const char* name, float* dst, size_t capacity, size_t* count,
                      int* shape, int* ndim);

/* HW1 (ref homeworks/hw1/src/template.c:195-238): C = A*B row block, init, check. */
void b200_cpu_matmul_f64(const double* A, const double* B, double* C, int rows, int n);
void b200_hw1_init_data(double* M, int rows, int cols);   /* fixed srand48(1) stream like init_data */
int  b200_hw1_check(const double* C, const double* Cref, size_t n, double tol);

/* ----------------------------------------------------------------- CUDA library ---- */

const char* b200_last_error(void);
int b200_cuda_device_count(void);

/* A row segment of a (possibly distributed) NHWC tensor: global rows [lo,hi) live in the slab
 * `ptr` (first slab row = global row `base`, `slab` rows per image).  `ptr` may be local memory
 * or a peer GPU's memory mapped over NVLink.  When `flag` is non-null the rows of image n may
 * only be read once flag[n] >= target (per-image arrival counter in local symmetric memory,
 * bumped by the producing rank's epilogue). */
typedef struct {
    const float* ptr;
    int base, slab;
    int lo, hi;
    const int* flag;
    int target;
    int pad_;
} b200_row_seg;

#define B200_MAX_SEG 4
#define B200_MAX_SIGNAL 4
#define B200_MAX_MIRROR 4

/* Halo push: pooled rows [lo,hi) of the block's output are ALSO stored into another slab (usually a
 * neighbour rank's copy of the activation, over NVLink), laid out like `out`: element (n,h,w,k) at
 * ((n*slab + h-base)*Wp + w)*K + k.  The reference sends these rows with MPI_Isend/Irecv after the layer
 * (ref v2_mpi_only/2.2_scatter_halo/src/main.cpp:178-187); here the producing epilogue writes them where the
 * consumer's loader will read them, so the consumer only ever reads local memory. */
typedef struct {
    float* ptr;
    int base, slab;
    int lo, hi;
} b200_row_mirror;

/* One fused block: conv(F,S,P) + bias + ReLU + max-pool(3,2) [+ cross-channel LRN], for pooled
 * rows [p0,p1) of every image.  Replaces the reference's 3-4 separate kernels per block
 * (ref v3_cuda_only/src/layers_cuda.cu:20-139, v4_mpi_cuda/src/layers_mpi_cuda.cu:25-89) and,
 * through `seg`/`signal`, its host-staged scatter / halo / gather
 * (ref v4_mpi_cuda/src/main_mpi_cuda.cpp:58-79,100-130). */
typedef struct {
    int N;                   /* images                                                    */
    int H, W, C;             /* logical conv input                                        */
    int K, F, S, P;
    int Hc, Wc;              /* logical conv output                                       */
    int Hp, Wp;              /* logical pooled output (pool window 3, stride 2)           */
    int p0, p1;              /* pooled rows to produce                                    */
    int out_base, out_slab;  /* output slab: (n,h,w,k) at ((n*out_slab+h-out_base)*Wp+w)*K+k */
    b200_lrn_params lrn;     /* N_lrn == 0 => no LRN                                      */
    int nseg;
    int n_signal;
    int in_pitch;            /* floats between consecutive input rows of every segment; 0 => W*C.  A
                                pitch that is a multiple of 4 lets conv1 (C=3) cp.async 16 B chunks  */
    b200_row_seg seg[B200_MAX_SEG];
    int* signal[B200_MAX_SIGNAL];   /* per-image counters (usually on peers) bumped per finished tile */
    int* status;             /* device word: first error code of the launch (0 = ok)      */
    unsigned long long timeout_ns;
    const int* epoch;        /* optional device counter of forwards: when set, seg[].target is a
                                per-forward increment and the wait target is (*epoch) * target, so
                                the launch arguments never change (CUDA-graph friendly)         */
    /* Flow control between forwards without a barrier (the role of the reference's MPI_Barrier pair,
     * ref v4_mpi_cuda/src/main_mpi_cuda.cpp:151,154): no CTA of the launch starts before *gate >=
     * gate_target (times *epoch when epoch is set); NULL = no gate.  The counter lives in local symmetric
     * memory and is bumped by the rank that owns the buffers this launch writes to. */
    const int* gate;
    int gate_target;
    /* Completion notice without an extra launch: every CTA adds 1 to *done_count (a local device
     * counter that therefore grows by the grid size per launch) after its last store, and the CTA that
     * completes a multiple of the grid size adds 1 to done[0..n_done) (usually rank 0's counter). */
    int n_done;
    unsigned int* done_count;
    int* done[2];
    /* signal[i] is bumped only by tiles that produce a pooled row in [signal_lo[i], signal_hi[i]) — the rows its
     * reader waits for — so interior tiles pay no system-scope release; hi <= lo means "every tile".  The
     * reader's target per image is b200_*_block_tiles_rows(g, lo, hi) per forward. */
    int signal_lo[B200_MAX_SIGNAL], signal_hi[B200_MAX_SIGNAL];
    int n_mirror;            /* halo push targets (see b200_row_mirror); stored before signal[] is bumped */
    int pad2_;
    b200_row_mirror mirror[B200_MAX_MIRROR];
} b200_block_geom;

/* host: [K][C][F][F] (reference layout) -> [F][F][C][K] used by the SIMT kernels */
int b200_pack_weights_ffck(const float* w_kcff, float* w_ffck, int K, int C, int F);

/* Exact-FP32 SIMT path (validation twin of the tensor-core path; V3 equivalent).
 * Number of tiles per image (needed by peers to compute flag targets) via *_tiles. */
int b200_simt_block(const float* w_ffck, const float* bias, float* out,
                    const b200_block_geom* g, void* stream);
int b200_simt_block_tiles(const b200_block_geom* g);
/* tiles per image that produce at least one pooled row of [lo,hi) (see signal_lo/signal_hi) */
int b200_simt_block_tiles_rows(const b200_block_geom* g, int lo, int hi);

/* tcgen05 / TMEM / TMA tensor-core path (TF32 operands, fp32 accumulate); same b200_block_geom.
 * A handle owns one layer's weights on the device, K-major [K][Kpad], TF32-rounded, plus the TMA
 * descriptor over them.  K must be a multiple of 32 and <= 256 (all channels in one CTA). */
int b200_tc_conv_create(const float* w_kcff, int K, int C, int F, void** handle);
int b200_tc_conv_destroy(void* handle);
int b200_tc_block(void* handle, const float* bias, float* out, const b200_block_geom* g,
                  int flags /* bit0: round outputs to TF32; bit1: programmatic dependent launch - the kernel may
                               become resident while the previous kernel in `stream` still runs and waits for its
                               completion before reading what it wrote (channel-split conv2 kernel; others ignore it) */,
                  void* stream);
int b200_tc_block_tiles(const b200_block_geom* g);
int b200_tc_block_tiles_rows(const b200_block_geom* g, int lo, int hi);
/* Process-wide kernel choice (default: environment B200_TC_CONV2 / B200_TC_CONV1 at first use, else auto):
 * key "conv2" = auto|single|db|pair|cs (single-buffered, double-buffered TMEM, tcgen05 CTA pair splitting the
 * pixels, CTA pair splitting the channels),
 * key "conv1" = auto|cpasync (pixel operand by TMA when eligible, or always the cp.async loader),
 * key "pdl" = 1|0 (honour flags bit1 of b200_tc_block; environment B200_TC_PDL),
 * key "dbg" = integer (timing experiments and plan-time variants of the conv kernels, see TcParams::dbg; 0 in production).
 * All choices produce bit-identical outputs; they differ in speed only (dbg bits below 1024 switch work off and
 * produce garbage). */
int b200_tc_set_policy(const char* key, const char* value);

/* Stand-alone convolution on NCHW device tensors with reference-layout weights [K][C][F][F]
 * (the `conv_test` equivalent; ref final_project/conv_test, kernel convTiledKernel).
 * out_nchw is [N][K][Ho][Wo], Ho = (H-F+2P)/S+1.  bias may be NULL.  The call allocates (and frees,
 * stream-ordered) b200_conv_nchw_scratch_bytes() of device scratch for its NHWC/packed intermediates. */
long long b200_conv_nchw_scratch_bytes(int N, int C, int H, int W, int K, int F, int S, int P);
int b200_conv_nchw(const float* in_nchw, const float* w_kcff, const float* bias, float* out_nchw,
                   int N, int C, int H, int W, int K, int F, int S, int P, void* stream);

/* Allocation-free variant of b200_conv_nchw.  It has no counterpart in the reference (whose CUDA
 * code allocates and frees on every call, ref v3_cuda_only/src/alexnet_cuda.cu:54-64,91-94).
 * For callers that must not allocate device memory when they call in: a server whose memory pool
 * already owns the whole GPU, or a latency-bound loop that cannot take cudaMalloc's device-wide
 * synchronisation.  Same arguments and results as b200_conv_nchw; the scratch comes from an 8 MiB
 * array built into the library (held by every GPU that loads it).  Problems needing more scratch
 * than b200_conv_nchw_noalloc_max_scratch_bytes() are refused with B200_ERR_RANGE.  All calls on one
 * GPU share that array: calls that overlap on the GPU race, so order them.  The first call on each
 * device resolves the array's address; issue it before any stream capture.  The array is part of the
 * library's device image, which the CUDA runtime sets up on a GPU at the first call into the library on
 * that GPU: make one warm-up call per device BEFORE a memory pool takes the rest of the GPU; from then on
 * the entry point neither allocates nor frees.  Not usable on a device after cudaDeviceReset(). */
long long b200_conv_nchw_noalloc_max_scratch_bytes(void);
int b200_conv_nchw_noalloc(const float* in_nchw, const float* w_kcff, const float* bias, float* out_nchw,
                           int N, int C, int H, int W, int K, int F, int S, int P, void* stream);

/* HW1 equivalent on the GPU: C[rows x n] = A[rows x n] * B[n x n], FP64 tensor cores (DMMA m8n8k4).
 * B and C may be peer-mapped pointers: the reference's MPI_Bcast(B) and MPI_Recv(C blocks)
 * (ref homeworks/hw1/src/template.c:132,141) become loads/stores inside the kernel. */
int b200_dgemm_rows(const double* A, const double* B, double* C, int rows, int n, void* stream);

/* ---- full-AlexNet extension (no counterpart in the reference beyond its plans,
 * ref final_project/PROBLEMS.txt:188-190, summary.md:37-43) ------------------------------------- */
/* tcgen05 TF32 GEMM: C[M][N] (+)= act(A[M][K] * W[N][K]^T + bias).  W is given on the host in
 * nn.Linear layout [N][K] (K % 32 == 0) and kept on the device TF32-rounded.  A and C may be peer
 * pointers (all-gather fused into the loader, reduce fused into the epilogue).
 * flags: 1 ReLU, 2 round outputs to TF32, 4 accumulate with red.add, 8 add bias. */
int b200_tc_gemm_create(const float* w_nk, int N, int K, void** handle);
int b200_tc_gemm_destroy(void* handle);
int b200_tc_gemm(void* handle, const float* A, int lda, int M, const float* bias, float* C, int ldc,
                 int flags, void* stream);
/* Split-K form for small-M layers: split ks of every tile covers a slice of K and stores its partial sums
 * at C + ks * split_stride floats (flags must be 0 when ksplit > 1); b200_splitk_reduce adds the partials in
 * a fixed order and applies bias / ReLU / TF32 rounding (flags 8 / 1 / 2).  b200_tc_gemm_info returns the
 * number of 32-wide k-blocks and the N tile (128 or 256) the kernel uses for this handle. */
int b200_tc_gemm_splitk(void* handle, const float* A, int lda, int M, const float* bias, float* C, int ldc,
                        int flags, int ksplit, long long split_stride, void* stream);
int b200_splitk_reduce(const float* ws, int ksplit, long long stride, long long rows, int cols,
                       const float* bias, float* out, int ld, int flags, void* stream);
int b200_tc_gemm_info(void* handle, int* nkb, int* bn);
/* Reduce-scatter by pull: out[r][c] = act(sum_k parts[k][r*ld_in + c] + bias[c]), k = 0..n-1 in order (n <= 16).
 * parts[k] may be peer-mapped pointers (every rank's partial products of the shard this rank owns): deterministic,
 * no atomics, no zeroed accumulator.  flags as b200_tc_gemm (1 ReLU, 2 round to TF32, 8 bias). */
int b200_peer_reduce(const float* const* parts, int n, long long rows, int cols, int ld_in, const float* bias,
                     float* out, int ld, int flags, void* stream);
/* Convolution as an implicit GEMM on the same kernel: x is NHWC [N][H][W][C] with C % 32 == 0 (may be a
 * peer pointer); the handle holds [K_out][F*F*C] weights in (fy, fx, c) order; out is [N*Ho*Wo][ldc].
 * The producer warps gather the im2col rows on the fly (zero fill outside the image); no col buffer. */
int b200_tc_conv_gemm(void* handle, const float* x, int N, int H, int W, int C, int F, int S, int P,
                      const float* bias, float* out, int ldc, int flags, void* stream);
int b200_im2col_nhwc(const float* in, float* col, int N, int H, int W, int C, int F, int S, int P, void* stream);
int b200_maxpool_nhwc(const float* in, float* out, int N, int H, int W, int C, int F, int S, void* stream);
int b200_bias_act(float* x, const float* bias, long long rows, int cols, int ld, int relu, void* stream);
int b200_softmax_rows(const float* x, float* y, int rows, int cols, int ld, void* stream);
/* contiguous rows of `width` floats -> rows `dpitch` floats apart (device to device) */
int b200_repitch_rows(const float* src, float* dst, long long rows, int width, int dpitch, void* stream);
/* space-to-depth by S of an NHWC row slab (rows [in_base,in_base+in_slab) of an H x W image) into s2d
 * rows [y0, y0+ys): out[n][Y-y0][X][(dy*S+dx)*C+c] = in[n][S*Y+dy][S*X+dx][c], zero outside. */
int b200_space_to_depth(const float* in, float* out, int N, int H, int W, int C, int S, int in_base,
                        int in_slab, int y0, int ys, void* stream);

/* Symmetric (peer-mapped) memory: zeroed cudaMalloc + CUDA IPC handles (64 bytes) exchanged by
 * the caller (torch.distributed); replaces MPI host buffers. */
int b200_symm_alloc(size_t bytes, void** ptr);
int b200_symm_free(void* ptr);
int b200_ipc_export(void* ptr, unsigned char* handle64);
int b200_ipc_open(const unsigned char* handle64, void** ptr);
int b200_ipc_close(void* ptr);
int b200_can_access_peer(int dev, int peer);
int b200_enable_peer(int peer);

/* Stream-ordered device flag ops (MPI_Barrier / completion notification equivalents,
 * ref v4_mpi_cuda/src/main_mpi_cuda.cpp:151,154): signal = fence.sys + red.release.sys add on up
 * to 16 counters (local or peer); wait = spin until flags[i] >= target[i] with a watchdog. */
/* pinned host -> device copy of `height` rows of `width` bytes with different pitches (cudaMemcpy2DAsync) */
int b200_memcpy2d_h2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t height,
                      void* stream);
/* device -> pinned host, rows of `width` bytes with different pitches: every rank writes the rows it owns
 * straight into one host tensor (the reference's D2H + MPI_Gatherv, ref v4_mpi_cuda/src/main_mpi_cuda.cpp:100-130) */
int b200_memcpy2d_d2h(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t height,
                      void* stream);
/* pin host memory this process did not allocate through CUDA (e.g. POSIX shm shared by all ranks) and return
 * its device-side address; b200_host_flag_store writes one int there in stream order (a "my copy has landed"
 * notice a reader in another process polls) */
int b200_host_register(void* host, size_t bytes, void** dev_ptr);
int b200_host_unregister(void* host);
int b200_host_flag_store(int* flag_dev_ptr, int value, void* stream);
int b200_flag_signal(int* const* targets, int n, int add, void* stream);
int b200_flag_wait(const int* const* flags, const int* target, int n, int* status,
                   unsigned long long timeout_ns, const int* epoch /* NULL or device counter: target *= *epoch */,
                   void* stream);
/* Opens a forward: ++*epoch (device counter of forwards, may be NULL) and adds 1 to each of the n (<= 16)
 * `go` counters (local or peer).  Rank 0 calls it with every rank's gate counter, the others with n = 0:
 * one tiny launch instead of an all-to-all barrier on the critical path. */
int b200_flag_begin(int* epoch, int* const* go_targets, int n, void* stream);
/* signal + wait in one launch (one GPU per rank only: the wait spins on peers' signals) */
int b200_flag_barrier(int* const* signal_targets, const int* const* wait_flags, const int* wait_target,
                      int n, int* status, unsigned long long timeout_ns,
                      int* epoch /* NULL, or device counter this kernel increments and waits for */,
                      void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200CNN_H */
