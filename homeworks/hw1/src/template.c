/* HW1 — distributed FP64 matrix multiply C = A*B, native: `mpirun -np N ./template <n>`.
 *
 * Same task and console contract as the course's program (ref homeworks/hw1/src/template.c:30-238): n a power of
 * two <= 4096 and divisible by the number of processes; the root owns A and B, every rank multiplies a block of
 * n/N rows of A by B, the root collects C, recomputes serially and prints "Test: PASSED|FAILED".
 * Here the row blocks go out with non-blocking sends while the root already works on its own block, B is
 * broadcast, and the arithmetic, the data generator and the check are the framework's CPU library
 * (b200_cpu_matmul_f64: cache-blocked, OpenMP over rows; b200_hw1_init_data; b200_hw1_check).
 * The GPU form of the same homework (row blocks on one GPU each, B read from and C written to the root's memory
 * over NVLink by an FP64 tensor-core kernel) is `python -m b200cnn.runners.hw1 --backend gpu` / `make gpu`.
 */
#include "b200cnn.h"

#include <mpi.h>
#include <stdio.h>
#include <stdlib.h>

#define ROOT 0
#define MAXDIM (1 << 12)

int main(int argc, char** argv) {
    int rank = 0, size = 1;
    MPI_Init(&argc, &argv);
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    MPI_Comm_size(MPI_COMM_WORLD, &size);
    int n = argc > 1 ? atoi(argv[1]) : 64;
    if (n <= 0 || (n & (n - 1)) != 0) {
        if (rank == ROOT) fprintf(stderr, "Error: Matrix dimension 'n' (%d) must be a positive power of two.\n", n);
        MPI_Abort(MPI_COMM_WORLD, 1);
    }
    if (n > MAXDIM) {
        if (rank == ROOT) fprintf(stderr, "Warning: Matrix dimension 'n' (%d) exceeds MAXDIM (%d). Clamping to MAXDIM.\n", n, MAXDIM);
        n = MAXDIM;
    }
    if (n % size != 0) {
        if (rank == ROOT)
            fprintf(stderr, "Error: Matrix dimension 'n' (%d) must be divisible by the number of processes (%d).\n", n, size);
        MPI_Abort(MPI_COMM_WORLD, 1);
    }
    const int my_work = n / size;
    const size_t nn = (size_t)n * n, blk = (size_t)my_work * n;
    double *A = NULL, *C = NULL;
    double* B = (double*)malloc(nn * sizeof(double));
    double* a_loc = (double*)malloc(blk * sizeof(double));
    double* c_loc = (double*)malloc(blk * sizeof(double));
    if (!B || !a_loc || !c_loc) MPI_Abort(MPI_COMM_WORLD, 2);
    if (rank == ROOT) {
        printf("pid=%d: num_procs=%d n=%d my_work=%d (rows per proc)\n", rank, size, n, my_work);
        A = (double*)malloc(nn * sizeof(double));
        C = (double*)malloc(nn * sizeof(double));
        if (!A || !C) MPI_Abort(MPI_COMM_WORLD, 2);
        b200_hw1_init_data(A, n, n);
        b200_hw1_init_data(B, n, n);
    }
    MPI_Barrier(MPI_COMM_WORLD);
    const double t0 = MPI_Wtime();
    MPI_Request* reqs = (MPI_Request*)malloc(sizeof(MPI_Request) * (size_t)size);
    if (rank == ROOT) {
        for (int p = 1; p < size; ++p)
            MPI_Isend(A + (size_t)p * blk, (int)blk, MPI_DOUBLE, p, 0, MPI_COMM_WORLD, &reqs[p]);
    } else {
        MPI_Recv(a_loc, (int)blk, MPI_DOUBLE, ROOT, 0, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
    }
    MPI_Bcast(B, (int)nn, MPI_DOUBLE, ROOT, MPI_COMM_WORLD);
    b200_cpu_matmul_f64(rank == ROOT ? A : a_loc, B, rank == ROOT ? C : c_loc, my_work, n);
    if (rank == ROOT) {
        if (size > 1) MPI_Waitall(size - 1, reqs + 1, MPI_STATUSES_IGNORE);
        for (int p = 1; p < size; ++p)
            MPI_Recv(C + (size_t)p * blk, (int)blk, MPI_DOUBLE, p, 1, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
    } else {
        MPI_Send(c_loc, (int)blk, MPI_DOUBLE, ROOT, 1, MPI_COMM_WORLD);
    }
    MPI_Barrier(MPI_COMM_WORLD);
    const double t1 = MPI_Wtime();
    int failed = 0;
    if (rank == ROOT) {
        printf("pid=%d: Parallel computation finished in %f seconds.\n", rank, t1 - t0);
        printf("pid=%d: Performing serial computation for verification...\n", rank);
        double* D = (double*)malloc(nn * sizeof(double));
        if (!D) MPI_Abort(MPI_COMM_WORLD, 2);
        const double s0 = MPI_Wtime();
        b200_cpu_matmul_f64(A, B, D, n, n);
        printf("pid=%d: Serial computation finished in %f seconds.\n", rank, MPI_Wtime() - s0);
        failed = b200_hw1_check(C, D, nn, 1e-6);
        printf("--------------------------------------\n");
        printf("pid=%d: Test: %s\n", rank, failed ? "FAILED" : "PASSED");
        if (!failed) printf("pid=%d: Total PARALLEL time: %f seconds.\n", rank, t1 - t0);
        printf("--------------------------------------\n");
        printf("backend=native-cpu\n");
        free(D);
    }
    free(reqs); free(A); free(C); free(B); free(a_loc); free(c_loc);
    MPI_Finalize();
    return failed ? 1 : 0;
}
