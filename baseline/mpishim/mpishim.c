/* Implementation of baseline/mpishim/mpi.h.  See the header for scope and transport. */
#define _GNU_SOURCE
#include "mpi.h"

#include <errno.h>
#include <fcntl.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <sys/types.h>
#include <unistd.h>

#ifdef B200_MPISHIM_CUDA
#include <cuda_runtime.h>
#include <dlfcn.h>
#endif

static int g_rank = 0, g_size = 1, g_init = 0;
static char g_dir[512];
static int* g_wfd;   /* g_wfd[d]: me -> d */
static int* g_rfd;   /* g_rfd[s]: s -> me */
static int* g_kfd;   /* g_kfd[s]: my own write end of s -> me, never written: keeps the fifo from reading EOF
                        while the real writer has not opened it yet (a rank that finished MPI_Init early
                        would otherwise see "no writer" = end of file in its first receive) */
static pthread_mutex_t* g_wlock;
static pthread_mutex_t* g_rlock;   /* per source: one reader at a time (send threads read their ACKs from the same pipe) */

typedef struct pend { int tag; int64_t n; char* data; struct pend* next; } pend;
static pend** g_pending;   /* per source: messages read while looking for another tag */

struct b200_mpi_req {
    int kind;            /* 0 send thread, 1 deferred recv */
    pthread_t th;
    void* buf; int64_t bytes; int peer, tag;
    void* owned;
};

static size_t dt_size(MPI_Datatype dt) {
    switch (dt) {
        case MPI_CHAR: case MPI_BYTE: return 1;
        case MPI_INT: case MPI_FLOAT: case MPI_UNSIGNED: return 4;
        case MPI_DOUBLE: case MPI_LONG: return 8;
        default: fprintf(stderr, "[mpishim] unknown datatype %d\n", dt); exit(70);
    }
}

static void die(const char* what) {
    fprintf(stderr, "[Rank %d] mpishim: %s: %s\n", g_rank, what, strerror(errno));
    exit(71);
}

static void full_write(int fd, const void* p, int64_t n) {
    const char* c = (const char*)p;
    while (n > 0) {
        ssize_t w = write(fd, c, (size_t)(n > (1 << 20) ? (1 << 20) : n));
        if (w < 0) { if (errno == EINTR) continue; die("write"); }
        c += w; n -= w;
    }
}
static void full_read(int fd, void* p, int64_t n) {
    char* c = (char*)p;
    while (n > 0) {
        ssize_t r = read(fd, c, (size_t)(n > (1 << 20) ? (1 << 20) : n));
        if (r < 0) { if (errno == EINTR) continue; die("read"); }
        if (r == 0) { fprintf(stderr, "[Rank %d] mpishim: peer closed the pipe\n", g_rank); exit(72); }
        c += r; n -= r;
    }
}

/* device-pointer aware staging ("CUDA-aware MPI" for the V5 reconstruction) */
static int is_device_ptr(const void* p) {
#ifdef B200_MPISHIM_CUDA
    struct cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return 0; }
    return a.type == cudaMemoryTypeDevice;
#else
    (void)p; return 0;
#endif
}
static void* stage_out(const void* buf, int64_t bytes) {   /* device -> host bounce */
#ifdef B200_MPISHIM_CUDA
    void* h = NULL;
    if (cudaMallocHost(&h, (size_t)bytes) != cudaSuccess) die("cudaMallocHost");
    cudaMemcpy(h, buf, (size_t)bytes, cudaMemcpyDeviceToHost);
    return h;
#else
    (void)buf; (void)bytes; return NULL;
#endif
}

/* ---- device-to-device transport ("CUDA-aware" path of the V5 reconstruction) --------------------------------
 * A device send buffer is not copied through the pipe: the sender ships a CUDA IPC handle of the allocation that
 * holds it (+ offset, length) and waits for an ACK; the receiver maps the allocation once (cached) and copies
 * straight from the sender's GPU into its own buffer with cudaMemcpy (peer copy over NVLink when the destination is
 * device memory).  The base of the allocation comes from the driver's cuMemGetAddressRange, looked up with dlopen so
 * the shim still loads on machines without a driver.  Anything that cannot take this path (no driver, host buffers)
 * falls back to the pinned bounce buffer above. */
#define TAG_ACK 0x7fff0003
static void raw_recv(void* buf, int64_t bytes, int src, int tag);
#ifdef B200_MPISHIM_CUDA
typedef struct { cudaIpcMemHandle_t h; uint64_t off; uint64_t bytes; } ipc_msg;
typedef int (*range_fn)(unsigned long long*, size_t*, unsigned long long);
static int dev_range(const void* p, void** base, size_t* size) {
    static range_fn fn = NULL; static int tried = 0;
    if (!tried) {
        tried = 1;
        void* lib = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
        if (lib) fn = (range_fn)dlsym(lib, "cuMemGetAddressRange_v2");
    }
    unsigned long long b = 0; size_t n = 0;
    if (!fn || fn(&b, &n, (unsigned long long)(uintptr_t)p) != 0) return 0;
    *base = (void*)(uintptr_t)b; *size = n; return 1;
}
static struct { cudaIpcMemHandle_t h; void* ptr; } g_ipc[64];
static int g_nipc = 0;
static pthread_mutex_t g_ipc_mu = PTHREAD_MUTEX_INITIALIZER;
static void* ipc_map(const cudaIpcMemHandle_t* h) {
    pthread_mutex_lock(&g_ipc_mu);
    for (int i = 0; i < g_nipc; ++i)
        if (!memcmp(&g_ipc[i].h, h, sizeof *h)) { void* p = g_ipc[i].ptr; pthread_mutex_unlock(&g_ipc_mu); return p; }
    void* p = NULL;
    if (cudaIpcOpenMemHandle(&p, *h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        fprintf(stderr, "[Rank %d] mpishim: cudaIpcOpenMemHandle failed: %s\n", g_rank, cudaGetErrorString(cudaGetLastError()));
        exit(74);
    }
    if (g_nipc < 64) { g_ipc[g_nipc].h = *h; g_ipc[g_nipc].ptr = p; ++g_nipc; }
    pthread_mutex_unlock(&g_ipc_mu);
    return p;
}
#endif

static void raw_send(const void* buf, int64_t bytes, int dest, int tag) {
    void* bounce = NULL;
#ifdef B200_MPISHIM_CUDA
    if (bytes > 0 && g_size > 1 && is_device_ptr(buf)) {
        void* base = NULL; size_t size = 0; ipc_msg m;
        if (dev_range(buf, &base, &size) && cudaIpcGetMemHandle(&m.h, base) == cudaSuccess) {
            m.off = (uint64_t)((const char*)buf - (const char*)base); m.bytes = (uint64_t)bytes;
            int32_t hdr_tag = tag; int64_t hdr_n = -(int64_t)sizeof m;       /* negative length = IPC descriptor */
            cudaDeviceSynchronize();                                         /* the data must be there before the peer reads it */
            pthread_mutex_lock(&g_wlock[dest]);
            full_write(g_wfd[dest], &hdr_tag, 4);
            full_write(g_wfd[dest], &hdr_n, 8);
            full_write(g_wfd[dest], &m, sizeof m);
            pthread_mutex_unlock(&g_wlock[dest]);
            raw_recv(NULL, 0, dest, TAG_ACK);                                /* buffer reusable once the peer has copied */
            return;
        }
        (void)cudaGetLastError();
    }
#endif
    if (bytes > 0 && is_device_ptr(buf)) { bounce = stage_out(buf, bytes); buf = bounce; }
    int32_t hdr_tag = tag; int64_t hdr_n = bytes;
    pthread_mutex_lock(&g_wlock[dest]);
    full_write(g_wfd[dest], &hdr_tag, 4);
    full_write(g_wfd[dest], &hdr_n, 8);
    full_write(g_wfd[dest], buf, bytes);
    pthread_mutex_unlock(&g_wlock[dest]);
#ifdef B200_MPISHIM_CUDA
    if (bounce) cudaFreeHost(bounce);
#endif
}

static void deliver(void* buf, int64_t want, const char* data, int64_t n) {
    if (n > want) { fprintf(stderr, "[Rank %d] mpishim: message truncated (%lld > %lld)\n", g_rank,
                            (long long)n, (long long)want); exit(73); }
#ifdef B200_MPISHIM_CUDA
    if (n > 0 && is_device_ptr(buf)) { cudaMemcpy(buf, data, (size_t)n, cudaMemcpyHostToDevice); return; }
#endif
    memcpy(buf, data, (size_t)n);
}

#define TAG_MATCH(want, got) ((want) == (got) || ((want) == MPI_ANY_TAG && (got) != TAG_ACK))

#ifdef B200_MPISHIM_CUDA
static void send_ack(int dest) {
    int32_t hdr_tag = TAG_ACK; int64_t hdr_n = 0;
    pthread_mutex_lock(&g_wlock[dest]);
    full_write(g_wfd[dest], &hdr_tag, 4);
    full_write(g_wfd[dest], &hdr_n, 8);
    pthread_mutex_unlock(&g_wlock[dest]);
}
#endif

static void raw_recv(void* buf, int64_t bytes, int src, int tag) {
    pthread_mutex_lock(&g_rlock[src]);
    /* 1. something already buffered from this source with this tag? */
    pend** pp = &g_pending[src];
    while (*pp) {
        if (TAG_MATCH(tag, (*pp)->tag)) {
            pend* m = *pp; *pp = m->next;
            deliver(buf, bytes, m->data, m->n);
            free(m->data); free(m);
            pthread_mutex_unlock(&g_rlock[src]);
            return;
        }
        pp = &(*pp)->next;
    }
    /* 2. read from the pipe until the tag shows up */
    for (;;) {
        int32_t t; int64_t n;
        full_read(g_rfd[src], &t, 4);
        full_read(g_rfd[src], &n, 8);
        char* data = NULL;
#ifdef B200_MPISHIM_CUDA
        if (n < 0) {                     /* IPC descriptor: the payload stays on the sender's GPU until we copy it */
            ipc_msg m;
            full_read(g_rfd[src], &m, sizeof m);
            const char* peer = (const char*)ipc_map(&m.h) + m.off;
            n = (int64_t)m.bytes;
            if (TAG_MATCH(tag, t)) {
                if (n > bytes) { fprintf(stderr, "[Rank %d] mpishim: message truncated (%lld > %lld)\n", g_rank,
                                         (long long)n, (long long)bytes); exit(73); }
                if (cudaMemcpy(buf, peer, (size_t)n, cudaMemcpyDefault) != cudaSuccess) die("cudaMemcpy (peer)");
                send_ack(src);
                pthread_mutex_unlock(&g_rlock[src]);
                return;
            }
            /* not the message we are looking for: take a host copy now, so its sender is released at once */
            data = (char*)malloc((size_t)(n > 0 ? n : 1));
            if (cudaMemcpy(data, peer, (size_t)n, cudaMemcpyDeviceToHost) != cudaSuccess) die("cudaMemcpy (peer->host)");
            send_ack(src);
        } else
#endif
        {
            data = (char*)malloc((size_t)(n > 0 ? n : 1));
            full_read(g_rfd[src], data, n);
            if (TAG_MATCH(tag, t)) { deliver(buf, bytes, data, n); free(data); pthread_mutex_unlock(&g_rlock[src]); return; }
        }
        pend* m = (pend*)malloc(sizeof(pend));
        m->tag = t; m->n = n; m->data = data; m->next = NULL;
        pend** q = &g_pending[src];
        while (*q) q = &(*q)->next;
        *q = m;
    }
}

static void fifo_path(char* out, size_t cap, int s, int d) { snprintf(out, cap, "%s/%d_%d", g_dir, s, d); }

int MPI_Init(int* argc, char*** argv) {
    (void)argc; (void)argv;
    if (g_init) return MPI_SUCCESS;
    const char* r = getenv("OMPI_COMM_WORLD_RANK"); const char* s = getenv("OMPI_COMM_WORLD_SIZE");
    g_rank = r ? atoi(r) : 0; g_size = s ? atoi(s) : 1;
    const char* d = getenv("B200_MPISHIM_DIR"); const char* port = getenv("MASTER_PORT");
    if (d) snprintf(g_dir, sizeof g_dir, "%s", d);
    else snprintf(g_dir, sizeof g_dir, "/tmp/b200mpi_%s_%d", port ? port : "solo", (int)getuid());
    g_wfd = (int*)calloc(g_size, sizeof(int)); g_rfd = (int*)calloc(g_size, sizeof(int));
    g_kfd = (int*)calloc(g_size, sizeof(int));
    g_wlock = (pthread_mutex_t*)calloc(g_size, sizeof(pthread_mutex_t));
    g_rlock = (pthread_mutex_t*)calloc(g_size, sizeof(pthread_mutex_t));
    for (int i = 0; i < g_size; ++i) pthread_mutex_init(&g_rlock[i], NULL);
    g_pending = (pend**)calloc(g_size, sizeof(pend*));
    for (int i = 0; i < g_size; ++i) { pthread_mutex_init(&g_wlock[i], NULL); g_wfd[i] = g_rfd[i] = g_kfd[i] = -1; }
    if (g_size > 1) {
        mkdir(g_dir, 0700);
        char p[600];
        for (int o = 0; o < g_size; ++o) {               /* create every fifo touching me */
            if (o == g_rank) continue;
            fifo_path(p, sizeof p, g_rank, o); if (mkfifo(p, 0600) && errno != EEXIST) die("mkfifo");
            fifo_path(p, sizeof p, o, g_rank); if (mkfifo(p, 0600) && errno != EEXIST) die("mkfifo");
        }
        /* open read ends non-blocking first (never blocks), then write ends (block until peer opened) */
        for (int o = 0; o < g_size; ++o) {
            if (o == g_rank) continue;
            fifo_path(p, sizeof p, o, g_rank);
            g_rfd[o] = open(p, O_RDONLY | O_NONBLOCK); if (g_rfd[o] < 0) die("open fifo (r)");
            g_kfd[o] = open(p, O_WRONLY | O_NONBLOCK); if (g_kfd[o] < 0) die("open fifo (keep-alive)");
        }
        for (int o = 0; o < g_size; ++o) {
            if (o == g_rank) continue;
            fifo_path(p, sizeof p, g_rank, o);
            g_wfd[o] = open(p, O_WRONLY); if (g_wfd[o] < 0) die("open fifo (w)");
            fcntl(g_wfd[o], F_SETPIPE_SZ, 1 << 20);
        }
        for (int o = 0; o < g_size; ++o)                  /* back to blocking reads */
            if (o != g_rank) fcntl(g_rfd[o], F_SETFL, fcntl(g_rfd[o], F_GETFL) & ~O_NONBLOCK);
    }
    g_init = 1;
    MPI_Barrier(MPI_COMM_WORLD);
    return MPI_SUCCESS;
}

int MPI_Finalize(void) {
    if (!g_init) return MPI_SUCCESS;
    MPI_Barrier(MPI_COMM_WORLD);
    char p[600];
    for (int o = 0; o < g_size; ++o) {
        if (o == g_rank) continue;
        close(g_wfd[o]); close(g_rfd[o]); close(g_kfd[o]);
        fifo_path(p, sizeof p, o, g_rank); unlink(p);
    }
    if (g_size > 1) rmdir(g_dir);
    g_init = 0;
    return MPI_SUCCESS;
}

int MPI_Comm_rank(MPI_Comm c, int* r) { (void)c; *r = g_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm c, int* s) { (void)c; *s = g_size; return MPI_SUCCESS; }
int MPI_Get_processor_name(char* n, int* l) { gethostname(n, MPI_MAX_PROCESSOR_NAME); *l = (int)strlen(n); return MPI_SUCCESS; }
double MPI_Wtime(void) { struct timeval t; gettimeofday(&t, NULL); return t.tv_sec + 1e-6 * t.tv_usec; }

int MPI_Abort(MPI_Comm c, int code) {
    (void)c;
    fprintf(stderr, "[Rank %d] MPI_Abort(%d)\n", g_rank, code);
    fflush(NULL);
    _exit(code ? code : 1);       /* bin/mpirun kills the other ranks (fail-stop) */
}

#define TAG_BARRIER 0x7fff0001
#define TAG_COLL 0x7fff0002

int MPI_Barrier(MPI_Comm c) {
    (void)c;
    if (g_size == 1) return MPI_SUCCESS;
    char z = 0;
    if (g_rank == 0) {
        for (int i = 1; i < g_size; ++i) raw_recv(&z, 1, i, TAG_BARRIER);
        for (int i = 1; i < g_size; ++i) raw_send(&z, 1, i, TAG_BARRIER);
    } else {
        raw_send(&z, 1, 0, TAG_BARRIER);
        raw_recv(&z, 1, 0, TAG_BARRIER);
    }
    return MPI_SUCCESS;
}

int MPI_Send(const void* b, int n, MPI_Datatype dt, int dest, int tag, MPI_Comm c) {
    (void)c; raw_send(b, (int64_t)n * dt_size(dt), dest, tag); return MPI_SUCCESS;
}
int MPI_Recv(void* b, int n, MPI_Datatype dt, int src, int tag, MPI_Comm c, MPI_Status* st) {
    (void)c; raw_recv(b, (int64_t)n * dt_size(dt), src, tag);
    if (st) { st->MPI_SOURCE = src; st->MPI_TAG = tag; st->MPI_ERROR = 0; }
    return MPI_SUCCESS;
}

static void* send_thread(void* a) {
    struct b200_mpi_req* r = (struct b200_mpi_req*)a;
    raw_send(r->buf, r->bytes, r->peer, r->tag);
    return NULL;
}
int MPI_Isend(const void* b, int n, MPI_Datatype dt, int dest, int tag, MPI_Comm c, MPI_Request* req) {
    (void)c;
    struct b200_mpi_req* r = (struct b200_mpi_req*)calloc(1, sizeof *r);
    r->kind = 0; r->buf = (void*)b; r->bytes = (int64_t)n * dt_size(dt); r->peer = dest; r->tag = tag;
    if (pthread_create(&r->th, NULL, send_thread, r)) die("pthread_create");
    *req = r; return MPI_SUCCESS;
}
int MPI_Irecv(void* b, int n, MPI_Datatype dt, int src, int tag, MPI_Comm c, MPI_Request* req) {
    (void)c;
    struct b200_mpi_req* r = (struct b200_mpi_req*)calloc(1, sizeof *r);
    r->kind = 1; r->buf = b; r->bytes = (int64_t)n * dt_size(dt); r->peer = src; r->tag = tag;
    *req = r; return MPI_SUCCESS;      /* the receive happens in Wait: sends never block on us */
}
int MPI_Wait(MPI_Request* req, MPI_Status* st) {
    struct b200_mpi_req* r = *req;
    if (!r) return MPI_SUCCESS;
    if (r->kind == 0) pthread_join(r->th, NULL);
    else raw_recv(r->buf, r->bytes, r->peer, r->tag);
    if (st) { st->MPI_SOURCE = r->peer; st->MPI_TAG = r->tag; st->MPI_ERROR = 0; }
    free(r); *req = MPI_REQUEST_NULL; return MPI_SUCCESS;
}
int MPI_Waitall(int n, MPI_Request* reqs, MPI_Status* sts) {
    for (int i = 0; i < n; ++i) if (reqs[i] && reqs[i]->kind == 1) MPI_Wait(&reqs[i], sts ? &sts[i] : NULL);
    for (int i = 0; i < n; ++i) if (reqs[i]) MPI_Wait(&reqs[i], sts ? &sts[i] : NULL);
    return MPI_SUCCESS;
}

int MPI_Bcast(void* b, int n, MPI_Datatype dt, int root, MPI_Comm c) {
    (void)c;
    int64_t bytes = (int64_t)n * dt_size(dt);
    if (g_size == 1) return MPI_SUCCESS;
    if (g_rank == root) { for (int i = 0; i < g_size; ++i) if (i != root) raw_send(b, bytes, i, TAG_COLL); }
    else raw_recv(b, bytes, root, TAG_COLL);
    return MPI_SUCCESS;
}

int MPI_Scatterv(const void* sb, const int* sc, const int* ds, MPI_Datatype sdt, void* rb, int rc,
                 MPI_Datatype rdt, int root, MPI_Comm c) {
    (void)c;
    if (g_rank == root) {
        size_t es = dt_size(sdt);
        for (int i = 0; i < g_size; ++i) {
            const char* src = (const char*)sb + (size_t)ds[i] * es;
            if (i == root) {
                if (sc[i] > 0) {
#ifdef B200_MPISHIM_CUDA
                    if (is_device_ptr(rb) || is_device_ptr(src)) cudaMemcpy(rb, src, (size_t)sc[i] * es, cudaMemcpyDefault);
                    else
#endif
                    memcpy(rb, src, (size_t)sc[i] * es);
                }
            } else raw_send(src, (int64_t)sc[i] * es, i, TAG_COLL);
        }
    } else raw_recv(rb, (int64_t)rc * dt_size(rdt), root, TAG_COLL);
    return MPI_SUCCESS;
}
int MPI_Scatter(const void* sb, int sc, MPI_Datatype sdt, void* rb, int rc, MPI_Datatype rdt, int root, MPI_Comm c) {
    int* cnt = (int*)malloc(sizeof(int) * g_size); int* dsp = (int*)malloc(sizeof(int) * g_size);
    for (int i = 0; i < g_size; ++i) { cnt[i] = sc; dsp[i] = i * sc; }
    int r = MPI_Scatterv(sb, cnt, dsp, sdt, rb, rc, rdt, root, c);
    free(cnt); free(dsp); return r;
}
int MPI_Gatherv(const void* sb, int sc, MPI_Datatype sdt, void* rb, const int* rcs, const int* ds,
                MPI_Datatype rdt, int root, MPI_Comm c) {
    (void)c;
    if (g_rank == root) {
        size_t es = dt_size(rdt);
        for (int i = 0; i < g_size; ++i) {
            char* dst = (char*)rb + (size_t)ds[i] * es;
            if (i == root) {
                if (sc > 0) {
#ifdef B200_MPISHIM_CUDA
                    if (is_device_ptr(dst) || is_device_ptr(sb)) cudaMemcpy(dst, sb, (size_t)sc * dt_size(sdt), cudaMemcpyDefault);
                    else
#endif
                    memcpy(dst, sb, (size_t)sc * dt_size(sdt));
                }
            } else raw_recv(dst, (int64_t)rcs[i] * es, i, TAG_COLL);
        }
    } else raw_send(sb, (int64_t)sc * dt_size(sdt), root, TAG_COLL);
    return MPI_SUCCESS;
}
int MPI_Gather(const void* sb, int sc, MPI_Datatype sdt, void* rb, int rc, MPI_Datatype rdt, int root, MPI_Comm c) {
    int* cnt = (int*)malloc(sizeof(int) * g_size); int* dsp = (int*)malloc(sizeof(int) * g_size);
    for (int i = 0; i < g_size; ++i) { cnt[i] = rc; dsp[i] = i * rc; }
    int r = MPI_Gatherv(sb, sc, sdt, rb, cnt, dsp, rdt, root, c);
    free(cnt); free(dsp); return r;
}
