// Native V2 `template`s: AlexNet blocks 1-2 on CPU ranks over MPI.
//   -DV2_BROADCAST_ALL   V2.1: rank 0 initialises, everything is broadcast, every rank computes the whole image,
//                        rank 0 prints (ref final_project/v2_mpi_only/2.1_broadcast_all/src/main.cpp:24-104)
//   (default)            V2.2: rows are scattered, pool1 halo rows exchanged between neighbours, final rows gathered
//                        (ref final_project/v2_mpi_only/2.2_scatter_halo/src/main.cpp:91-285)
// Unlike the reference's V2.2, which splits the INPUT rows and trims heuristically (and loses output rows at np > 1,
// SURVEY §0.2), the row ranges come from the exact planner (b200_plan_rank), so np N prints what np 1 prints.
// MPI surface: Init, rank/size, Bcast, Isend/Irecv/Waitall, Send/Recv, Barrier, Abort, Finalize — any MPI provides
// it; on this image `mpi.h` is the named-pipe subset in baseline/mpishim and the launcher is bin/mpirun.
// stdout contract (rank 0): "shape: HxWxC", "Sample values: ...", "Execution Time: X ms" (ref main.cpp:269-285).
#include "b200cnn.h"

#include <mpi.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

namespace {

struct Args {
    int H = 227, W = 227, C = 3, batch = 1;
    bool rand_init = false, extras = true;
    unsigned seed = 1;
    int lrn_mode = B200_LRN_ALPHA_OVER_N;
    int strategy = 0;        // 0 halo, 1 recompute
    const char* load_weights = nullptr;    // a weight file written by the Python side (utils/checkpoint.py)
};

const int K1 = 96, F1 = 11, S1 = 4, P1 = 0, K2 = 256, F2 = 5, S2 = 1, P2 = 2, FP = 3, SP = 2;

[[noreturn]] void die(int rank, const char* msg) {
    fprintf(stderr, "[Rank %d] %s\n", rank, msg);
    fflush(stderr);
    MPI_Abort(MPI_COMM_WORLD, 1);
    exit(1);
}

Args parse(int argc, char** argv, int rank) {
    Args a;
    for (int i = 1; i < argc; ++i) {
        auto val = [&](const char* flag) -> const char* {
            if (i + 1 >= argc) die(rank, (std::string("missing value after ") + flag).c_str());
            return argv[++i];
        };
        if (!strcmp(argv[i], "--height")) a.H = atoi(val("--height"));
        else if (!strcmp(argv[i], "--width")) a.W = atoi(val("--width"));
        else if (!strcmp(argv[i], "--batch")) a.batch = atoi(val("--batch"));
        else if (!strcmp(argv[i], "--seed")) a.seed = (unsigned)strtoul(val("--seed"), nullptr, 10);
        else if (!strcmp(argv[i], "--init")) a.rand_init = !strcmp(val("--init"), "rand");
        else if (!strcmp(argv[i], "--lrn")) {               // the Python runners' names: alpha_over_n (V1/V2), alpha_raw (V3/V4)
            const char* v = val("--lrn");
            a.lrn_mode = (!strcmp(v, "alpha_raw") || !strcmp(v, "v3")) ? B200_LRN_ALPHA_RAW : B200_LRN_ALPHA_OVER_N;
        }
        else if (!strcmp(argv[i], "--strategy")) a.strategy = !strcmp(val("--strategy"), "recompute") ? 1 : 0;
        else if (!strcmp(argv[i], "--quiet-extras")) a.extras = false;
        else if (!strcmp(argv[i], "--load-weights")) a.load_weights = val("--load-weights");
        else die(rank, (std::string("unknown argument ") + argv[i]).c_str());
    }
    if (a.H < F1 || a.W < F1 || a.batch < 1) die(rank, "bad --height / --width / --batch");
    return a;
}

struct Weights {
    std::vector<float> w1, b1, w2, b2;
};

// rank 0 draws input, W1, W2 in the reference's order; everybody then holds the same parameters
void init_problem(const Args& a, std::vector<float>& x, Weights& w) {
    x.resize((size_t)a.batch * a.H * a.W * a.C);
    w.w1.resize((size_t)K1 * a.C * F1 * F1);
    w.w2.resize((size_t)K2 * K1 * F2 * F2);
    if (a.rand_init) {
        b200_grand st;
        b200_grand_seed(&st, a.seed);
        b200_grand_fill_data(&st, x.data(), x.size());
        b200_grand_fill_weights(&st, w.w1.data(), w.w1.size());
        b200_grand_fill_weights(&st, w.w2.data(), w.w2.size());
        w.b1.assign(K1, 0.1f);
        w.b2.assign(K2, 0.1f);
    } else {                                   // ref main.cpp: input 1.0, weights 0.01, biases 0.0
        std::fill(x.begin(), x.end(), 1.0f);
        std::fill(w.w1.begin(), w.w1.end(), 0.01f);
        std::fill(w.w2.begin(), w.w2.end(), 0.01f);
        w.b1.assign(K1, 0.f);
        w.b2.assign(K2, 0.f);
    }
    if (a.load_weights) {                      // parameters from a file; the input stays as initialised above
        struct { const char* name; std::vector<float>* t; } want[] = {{"w1", &w.w1}, {"b1", &w.b1}, {"w2", &w.w2}, {"b2", &w.b2}};
        for (auto& q : want) {
            size_t n = 0;
            const int rc = b200_weights_read(a.load_weights, q.name, nullptr, 0, &n, nullptr, nullptr);
            if (rc || n != q.t->size() || b200_weights_read(a.load_weights, q.name, q.t->data(), q.t->size(), &n, nullptr, nullptr))
                die(0, (std::string("--load-weights ") + a.load_weights + ": cannot read " + q.name +
                        " (missing file, wrong shape or checksum mismatch)").c_str());
        }
    }
}

void bcast_weights(const Args& a, Weights& w, int rank) {
    if (rank != 0) {
        w.w1.resize((size_t)K1 * a.C * F1 * F1);
        w.b1.resize(K1);
        w.w2.resize((size_t)K2 * K1 * F2 * F2);
        w.b2.resize(K2);
    }
    MPI_Bcast(w.w1.data(), (int)w.w1.size(), MPI_FLOAT, 0, MPI_COMM_WORLD);
    MPI_Bcast(w.b1.data(), (int)w.b1.size(), MPI_FLOAT, 0, MPI_COMM_WORLD);
    MPI_Bcast(w.w2.data(), (int)w.w2.size(), MPI_FLOAT, 0, MPI_COMM_WORLD);
    MPI_Bcast(w.b2.data(), (int)w.b2.size(), MPI_FLOAT, 0, MPI_COMM_WORLD);
}

#ifndef V2_BROADCAST_ALL
// rows [lo,hi) of every image of a slab tensor [N][slab rows from `base`][rowlen] -> dense [N][hi-lo][rowlen]
std::vector<float> pack_rows(const std::vector<float>& t, int N, int base, int slab, int lo, int hi, size_t rowlen) {
    std::vector<float> out((size_t)N * (hi - lo) * rowlen);
    for (int n = 0; n < N; ++n)
        memcpy(out.data() + (size_t)n * (hi - lo) * rowlen, t.data() + ((size_t)n * slab + (lo - base)) * rowlen,
               (size_t)(hi - lo) * rowlen * sizeof(float));
    return out;
}

void unpack_rows(std::vector<float>& t, int N, int base, int slab, int lo, int hi, size_t rowlen,
                 const std::vector<float>& src) {
    for (int n = 0; n < N; ++n)
        memcpy(t.data() + ((size_t)n * slab + (lo - base)) * rowlen, src.data() + (size_t)n * (hi - lo) * rowlen,
               (size_t)(hi - lo) * rowlen * sizeof(float));
}


enum { TAG_SCATTER = 10, TAG_HALO = 2, TAG_GATHER = 20 };   // the reference uses 0/1 (halo 1) and 2/3 (halo 2)
#endif

// conv1 -> ReLU -> pool1 for pool1 rows [own0,own1) from the input slab [in0,in1)
std::vector<float> block1(const Args& a, const Weights& w, const b200_rank_plan& p, const std::vector<float>& in,
                          int Wc1, int Wp1, int rank) {
    const int nc = p.conv1_1 - p.conv1_0, no = p.own1 - p.own0;
    std::vector<float> c((size_t)a.batch * nc * Wc1 * K1), o((size_t)a.batch * no * Wp1 * K1);
    if (no <= 0) return o;
    b200_win_geom g = {a.batch, a.H, a.W, a.C, p.in0, p.in1 - p.in0, K1, F1, S1, P1, p.Hc1, Wc1,
                       p.conv1_0, nc, p.conv1_0, p.conv1_1};
    if (b200_cpu_conv(in.data(), w.w1.data(), w.b1.data(), c.data(), &g)) die(rank, "conv1 failed");
    b200_cpu_relu(c.data(), c.size());
    g = {a.batch, p.Hc1, Wc1, K1, p.conv1_0, nc, K1, FP, SP, 0, p.Hp1, Wp1, p.own0, no, p.own0, p.own1};
    if (b200_cpu_maxpool(c.data(), o.data(), &g)) die(rank, "pool1 failed");
    return o;
}

// conv2 -> ReLU -> pool2 -> LRN2 for final rows [out0,out1) from the pool1 slab that holds [need0,need1)
std::vector<float> block2(const Args& a, const Weights& w, const b200_rank_plan& p, const std::vector<float>& slab,
                          int Wp1, int Wc2, int Wp2, int rank) {
    const int nc = p.conv2_1 - p.conv2_0, no = p.out1 - p.out0;
    std::vector<float> c((size_t)a.batch * nc * Wc2 * K2), o((size_t)a.batch * no * Wp2 * K2), y(o.size());
    if (no <= 0) return y;
    b200_win_geom g = {a.batch, p.Hp1, Wp1, K1, p.need0, p.need1 - p.need0, K2, F2, S2, P2, p.Hc2, Wc2,
                       p.conv2_0, nc, p.conv2_0, p.conv2_1};
    if (b200_cpu_conv(slab.data(), w.w2.data(), w.b2.data(), c.data(), &g)) die(rank, "conv2 failed");
    b200_cpu_relu(c.data(), c.size());
    g = {a.batch, p.Hc2, Wc2, K2, p.conv2_0, nc, K2, FP, SP, 0, p.Hp2, Wp2, p.out0, no, p.out0, p.out1};
    if (b200_cpu_maxpool(c.data(), o.data(), &g)) die(rank, "pool2 failed");
    b200_lrn_params lp = {5, 0.0001f, 0.75f, 2.0f, a.lrn_mode};
    if (b200_cpu_lrn(o.data(), y.data(), (size_t)a.batch * no * Wp2, K2, &lp)) die(rank, "lrn2 failed");
    return y;
}

}  // namespace

int main(int argc, char** argv) {
    MPI_Init(&argc, &argv);
    int rank = 0, world = 1;
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    MPI_Comm_size(MPI_COMM_WORLD, &world);
    const Args a = parse(argc, argv, rank);
    const b200_plan_block pb1 = {F1, S1, P1, FP, SP}, pb2 = {F2, S2, P2, FP, SP};
    b200_rank_plan whole;
    if (b200_plan_rank(a.H, a.batch, &pb1, &pb2, 1, 0, 0, 0, &whole)) die(rank, "planner failed");
    const int Wc1 = (a.W - F1 + 2 * P1) / S1 + 1, Wp1 = (Wc1 - FP) / SP + 1;
    const int Wc2 = (Wp1 - F2 + 2 * P2) / S2 + 1, Wp2 = (Wc2 - FP) / SP + 1;
    if (whole.Hp2 <= 0 || Wp2 <= 0) die(rank, "input too small for blocks 1-2");

    std::vector<float> x;
    Weights w;
    if (rank == 0) init_problem(a, x, w);
    bcast_weights(a, w, rank);
    std::vector<float> y;                        // rank 0: [N][Hp2][Wp2][K2]

#ifdef V2_BROADCAST_ALL
    x.resize((size_t)a.batch * a.H * a.W * a.C);
    MPI_Bcast(x.data(), (int)x.size(), MPI_FLOAT, 0, MPI_COMM_WORLD);
    MPI_Barrier(MPI_COMM_WORLD);
    const auto t0 = std::chrono::high_resolution_clock::now();
    {
        const std::vector<float> p1 = block1(a, w, whole, x, Wc1, Wp1, rank);
        y = block2(a, w, whole, p1, Wp1, Wc2, Wp2, rank);
    }
    MPI_Barrier(MPI_COMM_WORLD);
#else
    std::vector<b200_rank_plan> plans(world);
    for (int r = 0; r < world; ++r)
        if (b200_plan_rank(a.H, a.batch, &pb1, &pb2, world, a.strategy, 0, r, &plans[r])) die(rank, "planner failed");
    const b200_rank_plan& me = plans[rank];
    MPI_Barrier(MPI_COMM_WORLD);
    const auto t0 = std::chrono::high_resolution_clock::now();
    // -- scatter: every rank gets exactly the input rows its conv1 rows need ------------------------------------
    const size_t in_row = (size_t)a.W * a.C;
    std::vector<float> mine;
    if (rank == 0) {
        std::vector<std::vector<float>> bufs(world);
        std::vector<MPI_Request> reqs;
        for (int r = 1; r < world; ++r) {
            if (plans[r].in1 <= plans[r].in0) continue;
            bufs[r] = pack_rows(x, a.batch, 0, a.H, plans[r].in0, plans[r].in1, in_row);
            reqs.emplace_back();
            MPI_Isend(bufs[r].data(), (int)bufs[r].size(), MPI_FLOAT, r, TAG_SCATTER, MPI_COMM_WORLD, &reqs.back());
        }
        mine = pack_rows(x, a.batch, 0, a.H, me.in0, me.in1, in_row);
        if (!reqs.empty()) MPI_Waitall((int)reqs.size(), reqs.data(), MPI_STATUSES_IGNORE);
    } else {
        mine.resize((size_t)a.batch * (me.in1 - me.in0) * in_row);
        if (me.in1 > me.in0)
            MPI_Recv(mine.data(), (int)mine.size(), MPI_FLOAT, 0, TAG_SCATTER, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
    }
    // -- block 1 on the owned pool1 rows, then the halo exchange (post receives, then sends, then wait) ---------
    const std::vector<float> own = block1(a, w, me, mine, Wc1, Wp1, rank);
    const size_t p1_row = (size_t)Wp1 * K1;
    const int nneed = me.need1 - me.need0;
    std::vector<float> slab((size_t)a.batch * std::max(nneed, 0) * p1_row, 0.f);
    {
        const int lo = std::max(me.own0, me.need0), hi = std::min(me.own1, me.need1);
        if (hi > lo) unpack_rows(slab, a.batch, me.need0, nneed, lo, hi, p1_row,
                                 pack_rows(own, a.batch, me.own0, me.own1 - me.own0, lo, hi, p1_row));
        std::vector<std::vector<float>> rbuf(me.n_halo), sbuf(me.n_serves);
        std::vector<MPI_Request> reqs(me.n_halo + me.n_serves);
        for (int i = 0; i < me.n_halo; ++i) {
            rbuf[i].resize((size_t)a.batch * (me.halo[i].hi - me.halo[i].lo) * p1_row);
            MPI_Irecv(rbuf[i].data(), (int)rbuf[i].size(), MPI_FLOAT, me.halo[i].peer, TAG_HALO, MPI_COMM_WORLD, &reqs[i]);
        }
        for (int i = 0; i < me.n_serves; ++i) {
            sbuf[i] = pack_rows(own, a.batch, me.own0, me.own1 - me.own0, me.serves[i].lo, me.serves[i].hi, p1_row);
            MPI_Isend(sbuf[i].data(), (int)sbuf[i].size(), MPI_FLOAT, me.serves[i].peer, TAG_HALO, MPI_COMM_WORLD,
                      &reqs[me.n_halo + i]);
        }
        if (!reqs.empty()) MPI_Waitall((int)reqs.size(), reqs.data(), MPI_STATUSES_IGNORE);
        for (int i = 0; i < me.n_halo; ++i)
            unpack_rows(slab, a.batch, me.need0, nneed, me.halo[i].lo, me.halo[i].hi, p1_row, rbuf[i]);
    }
    // -- block 2 on the owned final rows; rank 0 collects the rows at their place (no trim: ownership is exact) --
    const std::vector<float> part = block2(a, w, me, slab, Wp1, Wc2, Wp2, rank);
    const size_t out_row = (size_t)Wp2 * K2;
    if (rank == 0) {
        y.assign((size_t)a.batch * whole.Hp2 * out_row, 0.f);
        if (me.out1 > me.out0) unpack_rows(y, a.batch, 0, whole.Hp2, me.out0, me.out1, out_row, part);
        for (int r = 1; r < world; ++r) {
            if (plans[r].out1 <= plans[r].out0) continue;
            std::vector<float> buf((size_t)a.batch * (plans[r].out1 - plans[r].out0) * out_row);
            MPI_Recv(buf.data(), (int)buf.size(), MPI_FLOAT, r, TAG_GATHER, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
            unpack_rows(y, a.batch, 0, whole.Hp2, plans[r].out0, plans[r].out1, out_row, buf);
        }
    } else if (me.out1 > me.out0) {
        MPI_Send(part.data(), (int)part.size(), MPI_FLOAT, 0, TAG_GATHER, MPI_COMM_WORLD);
    }
    MPI_Barrier(MPI_COMM_WORLD);
#endif
    const auto t1 = std::chrono::high_resolution_clock::now();
    if (rank == 0) {
        const double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
        std::cout << "shape: " << whole.Hp2 << "x" << Wp2 << "x" << K2 << std::endl;
        std::cout << "Sample values:";
        for (int i = 0; i < 5 && i < (int)y.size(); ++i) std::cout << " " << y[i];
        std::cout << std::endl;
        std::cout << "Execution Time: " << ms << " ms" << std::endl;
        if (a.extras) {
            printf("Output FNV1a64: 0x%016llx\n", (unsigned long long)b200_fnv1a64(y.data(), y.size() * 4));
            printf("np=%d\nbatch=%d\nnative=1\n", world, a.batch);
            printf("ExecutionTime_ms: %g\n", ms);
        }
    }
    MPI_Finalize();
    return 0;
}
