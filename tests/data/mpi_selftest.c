/* Exercises every call of baseline/mpishim/mpi.h (the subset the reference uses). */
#include <mpi.h>
#include <stdio.h>
#include <stdlib.h>

int main(int argc, char** argv) {
    MPI_Init(&argc, &argv);
    int rank, size, ok = 1;
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    MPI_Comm_size(MPI_COMM_WORLD, &size);
    /* Bcast */
    double b[3] = {0, 0, 0};
    if (rank == 0) { b[0] = 1.5; b[1] = 2.5; b[2] = 3.5; }
    MPI_Bcast(b, 3, MPI_DOUBLE, 0, MPI_COMM_WORLD);
    ok &= b[2] == 3.5;
    /* Scatterv uneven rows, ring halo with Isend/Irecv (tags per direction), Gatherv back */
    int* cnt = malloc(sizeof(int) * size); int* disp = malloc(sizeof(int) * size);
    int total = 0;
    for (int r = 0; r < size; ++r) { cnt[r] = 2 + r; disp[r] = total; total += cnt[r]; }
    float* full = NULL;
    if (rank == 0) { full = malloc(sizeof(float) * total); for (int i = 0; i < total; ++i) full[i] = (float)i; }
    float* mine = malloc(sizeof(float) * cnt[rank]);
    MPI_Scatterv(full, cnt, disp, MPI_FLOAT, mine, cnt[rank], MPI_FLOAT, 0, MPI_COMM_WORLD);
    ok &= mine[0] == (float)disp[rank];
    float up = -1, down = -1; MPI_Request rq[4]; int q = 0;
    if (rank > 0) MPI_Irecv(&up, 1, MPI_FLOAT, rank - 1, 0, MPI_COMM_WORLD, &rq[q++]);
    if (rank < size - 1) MPI_Irecv(&down, 1, MPI_FLOAT, rank + 1, 1, MPI_COMM_WORLD, &rq[q++]);
    if (rank > 0) MPI_Isend(&mine[0], 1, MPI_FLOAT, rank - 1, 1, MPI_COMM_WORLD, &rq[q++]);
    if (rank < size - 1) MPI_Isend(&mine[cnt[rank] - 1], 1, MPI_FLOAT, rank + 1, 0, MPI_COMM_WORLD, &rq[q++]);
    MPI_Waitall(q, rq, MPI_STATUSES_IGNORE);
    if (rank > 0) ok &= up == (float)(disp[rank] - 1);
    if (rank < size - 1) ok &= down == (float)(disp[rank] + cnt[rank]);
    for (int i = 0; i < cnt[rank]; ++i) mine[i] *= 2;
    MPI_Gatherv(mine, cnt[rank], MPI_FLOAT, full, cnt, disp, MPI_FLOAT, 0, MPI_COMM_WORLD);
    if (rank == 0) for (int i = 0; i < total; ++i) ok &= full[i] == 2.0f * i;
    /* Send/Recv with out-of-order tags, Scatter/Gather of ints, Barrier, Wtime */
    if (size > 1) {
        int a = 11, c = 22, ra = 0, rc = 0;
        if (rank == 0) { MPI_Send(&a, 1, MPI_INT, 1, 7, MPI_COMM_WORLD); MPI_Send(&c, 1, MPI_INT, 1, 8, MPI_COMM_WORLD); }
        if (rank == 1) { MPI_Recv(&rc, 1, MPI_INT, 0, 8, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
                         MPI_Recv(&ra, 1, MPI_INT, 0, 7, MPI_COMM_WORLD, MPI_STATUS_IGNORE); ok &= ra == 11 && rc == 22; }
    }
    int mynum = -1; int* nums = malloc(sizeof(int) * size);
    for (int r = 0; r < size; ++r) nums[r] = 100 + r;
    MPI_Scatter(nums, 1, MPI_INT, &mynum, 1, MPI_INT, 0, MPI_COMM_WORLD);
    ok &= mynum == 100 + rank;
    mynum += 1;
    MPI_Gather(&mynum, 1, MPI_INT, nums, 1, MPI_INT, 0, MPI_COMM_WORLD);
    if (rank == 0) for (int r = 0; r < size; ++r) ok &= nums[r] == 101 + r;
    MPI_Barrier(MPI_COMM_WORLD);
    double t = MPI_Wtime(); ok &= t > 0;
    int allok = ok;
    /* everyone reports to rank 0 */
    if (rank == 0) { for (int r = 1; r < size; ++r) { int o; MPI_Recv(&o, 1, MPI_INT, r, 99, MPI_COMM_WORLD, MPI_STATUS_IGNORE); allok &= o; }
                     printf("mpishim selftest np=%d: %s\n", size, allok ? "PASSED" : "FAILED"); }
    else MPI_Send(&ok, 1, MPI_INT, 0, 99, MPI_COMM_WORLD);
    free(cnt); free(disp); free(full); free(mine); free(nums);
    MPI_Finalize();
    return allok ? 0 : 1;
}
