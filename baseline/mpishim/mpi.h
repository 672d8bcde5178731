/* mpi.h — the 18-call MPI subset the reference uses, implemented over named pipes.
 *
 * The image ships no MPI.  Compiled against this header and launched with bin/mpirun (which provides
 * OMPI_COMM_WORLD_RANK/SIZE): the UNMODIFIED reference sources (V2.1, V2.2, V4, HW1) as same-box baselines, and
 * the framework's own native MPI programs — the V2 `template` binaries (csrc/apps/v2_main.cpp) and the native
 * homework skeleton (templates/template.c.template).  With a real MPI installed they build against that instead
 * (make CXX=mpicxx MPI_INC= MPI_SRC=); the GPU paths never use MPI at all (peer memory + torch.distributed bootstrap).
 *
 * Transport: one FIFO per ordered rank pair under $B200_MPISHIM_DIR (default /tmp/b200mpi_<port>);
 * messages are [tag,int32][bytes,int64][payload]; Isend runs on a helper thread so symmetric
 * Isend/Irecv exchanges cannot deadlock; collectives are linear trees on top of Send/Recv.
 * MPI_COMM_WORLD only; MPI_INT / MPI_FLOAT / MPI_DOUBLE / MPI_CHAR / MPI_BYTE.
 *
 * "V5 reconstruction" (SURVEY §0.2/§6.3): when a buffer passed to Send/Recv/Isend/Irecv/Scatterv/
 * Gatherv is a CUDA device pointer and the shim was built with -DB200_MPISHIM_CUDA, it is staged
 * through a pinned bounce buffer transparently — i.e. a CUDA-aware MPI, which is what the
 * reference's V5 was specified to rely on (README.md:158-166).
 */
#ifndef B200_MPISHIM_MPI_H
#define B200_MPISHIM_MPI_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;
typedef struct b200_mpi_req* MPI_Request;

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_CHAR 1
#define MPI_BYTE 2
#define MPI_INT 3
#define MPI_FLOAT 4
#define MPI_DOUBLE 5
#define MPI_UNSIGNED 6
#define MPI_LONG 7
#define MPI_ANY_TAG (-1)
#define MPI_ANY_SOURCE (-2)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
#define MPI_REQUEST_NULL ((MPI_Request)0)
#define MPI_MAX_PROCESSOR_NAME 256

int MPI_Init(int* argc, char*** argv);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Barrier(MPI_Comm comm);
int MPI_Abort(MPI_Comm comm, int code);
double MPI_Wtime(void);
int MPI_Get_processor_name(char* name, int* len);

int MPI_Send(const void* buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void* buf, int count, MPI_Datatype dt, int src, int tag, MPI_Comm comm, MPI_Status* st);
int MPI_Isend(const void* buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request* req);
int MPI_Irecv(void* buf, int count, MPI_Datatype dt, int src, int tag, MPI_Comm comm, MPI_Request* req);
int MPI_Wait(MPI_Request* req, MPI_Status* st);
int MPI_Waitall(int n, MPI_Request* reqs, MPI_Status* sts);

int MPI_Bcast(void* buf, int count, MPI_Datatype dt, int root, MPI_Comm comm);
int MPI_Scatter(const void* sbuf, int scount, MPI_Datatype sdt, void* rbuf, int rcount, MPI_Datatype rdt,
                int root, MPI_Comm comm);
int MPI_Scatterv(const void* sbuf, const int* scounts, const int* displs, MPI_Datatype sdt, void* rbuf,
                 int rcount, MPI_Datatype rdt, int root, MPI_Comm comm);
int MPI_Gather(const void* sbuf, int scount, MPI_Datatype sdt, void* rbuf, int rcount, MPI_Datatype rdt,
               int root, MPI_Comm comm);
int MPI_Gatherv(const void* sbuf, int scount, MPI_Datatype sdt, void* rbuf, const int* rcounts,
                const int* displs, MPI_Datatype rdt, int root, MPI_Comm comm);

#ifdef __cplusplus
}
#endif
#endif
