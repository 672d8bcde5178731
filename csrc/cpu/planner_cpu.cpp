// Exact row (x batch) decomposition planner for conv -> pool -> conv -> pool stacks, native twin of
// cuda-mpi-gpu-cluster-programming_b200/parallel/planner.py (the two are compared by tests/test_planner.py).
//
// The reference splits the INPUT rows (base = H/size, the first H % size ranks get one more), exchanges F/2 halo
// rows and trims heuristically (ref final_project/v2_mpi_only/2.2_scatter_halo/src/main.cpp:101-116,118-135,159-176;
// final_project/v4_mpi_cuda/src/main_mpi_cuda.cpp:55-79,102-122) and loses output rows doing so.  Here
// the FINAL rows are partitioned with the same base+remainder rule and every earlier layer's range follows from the
// receptive-field arithmetic, so the union of the ranks' outputs is exactly the one-rank output.  A native host (the
// V2 `template` binaries, a C++ serving process driving the C ABI) gets its slab geometry from here.
#include "b200cnn.h"

#include <algorithm>
#include <vector>

namespace {

struct Range { int lo, hi; };

inline int len(Range r) { return std::max(0, r.hi - r.lo); }

inline int conv_out(int D, int F, int P, int S) { return D + 2 * P < F ? 0 : (D - F + 2 * P) / S + 1; }
inline int pool_out(int D, int F, int S) { return D < F ? 0 : (D - F) / S + 1; }

// base = n / parts rows each, the first n % parts get one more (ref v2.2 main.cpp:102-109)
std::vector<Range> split_rows(int n, int parts) {
    std::vector<Range> out;
    const int base = n / parts, rem = n % parts;
    int off = 0;
    for (int r = 0; r < parts; ++r) {
        const int k = base + (r < rem ? 1 : 0);
        out.push_back({off, off + k});
        off += k;
    }
    return out;
}

// input rows touched by output rows `o` of a window op, clipped to the tensor
Range window_in(Range o, int F, int S, int P, int D) {
    if (o.hi <= o.lo) return {0, 0};
    return {std::max(o.lo * S - P, 0), std::min((o.hi - 1) * S - P + F, D)};
}

Range intersect(Range a, Range b) {
    const int lo = std::max(a.lo, b.lo), hi = std::min(a.hi, b.hi);
    return hi > lo ? Range{lo, hi} : Range{lo, lo};
}

struct RowPlan {
    Range out, conv2, need, own, conv1, in;
    std::vector<b200_plan_halo> halo, serves;
};

}  // namespace

extern "C" int b200_plan_rank(int H, int batch, const b200_plan_block* b1, const b200_plan_block* b2, int world,
                              int strategy, int row_groups, int rank, b200_rank_plan* out) {
    if (!b1 || !b2 || !out || H <= 0 || batch < 0 || world < 1 || rank < 0 || rank >= world) return B200_ERR_ARG;
    if (b1->S < 1 || b2->S < 1 || b1->S_pool < 1 || b2->S_pool < 1 || (strategy != 0 && strategy != 1))
        return B200_ERR_ARG;
    const int R = row_groups > 0 ? row_groups : world;
    if (world % R != 0) return B200_ERR_ARG;
    const int B = world / R;
    const int Hc1 = conv_out(H, b1->F, b1->P, b1->S), Hp1 = pool_out(Hc1, b1->F_pool, b1->S_pool);
    const int Hc2 = conv_out(Hp1, b2->F, b2->P, b2->S), Hp2 = pool_out(Hc2, b2->F_pool, b2->S_pool);

    const std::vector<Range> finals = split_rows(Hp2, R);
    // pool1 ownership: pool1 row p feeds conv2 row p (stride-1 centre), which falls into pool2 row
    // floor(c / S_pool) (clipped) -> the owner of that final row.  Monotone, so owned rows are contiguous.
    std::vector<Range> own(R, Range{-1, -1});
    if (Hp2 > 0) {
        std::vector<int> row_owner(Hp2, 0);
        for (int r = 0; r < R; ++r)
            for (int f = finals[r].lo; f < finals[r].hi; ++f) row_owner[f] = r;
        for (int p = 0; p < Hp1; ++p) {
            int c = Hc2 ? std::min(std::max((p + b2->P - b2->F / 2) / std::max(b2->S, 1), 0), Hc2 - 1) : 0;
            const int f = std::min(c / b2->S_pool, Hp2 - 1);
            const int r = row_owner[f];
            if (own[r].lo < 0) own[r].lo = p;
            own[r].hi = p + 1;
        }
    }
    std::vector<RowPlan> g(R);
    for (int r = 0; r < R; ++r) {
        RowPlan& q = g[r];
        q.out = finals[r];
        q.conv2 = window_in(q.out, b2->F_pool, b2->S_pool, 0, Hc2);
        q.need = window_in(q.conv2, b2->F, b2->S, b2->P, Hp1);
        if (strategy == 0) q.own = own[r].lo >= 0 ? own[r] : Range{q.need.lo, q.need.lo};
        else q.own = q.need;
        q.conv1 = window_in(q.own, b1->F_pool, b1->S_pool, 0, Hc1);
        q.in = window_in(q.conv1, b1->F, b1->S, b1->P, H);
    }
    if (strategy == 0)
        for (int r = 0; r < R; ++r) {
            if (len(g[r].out) == 0) continue;
            for (int o = 0; o < R; ++o) {
                if (o == r || len(g[o].own) == 0) continue;
                const Range ov = intersect(g[r].need, g[o].own);
                if (len(ov)) {
                    g[r].halo.push_back({o, ov.lo, ov.hi});
                    g[o].serves.push_back({r, ov.lo, ov.hi});
                }
            }
        }
    const int q = rank % R, bg = rank / R;
    const RowPlan& me = g[q];
    if ((int)me.halo.size() > B200_PLAN_MAX_HALO || (int)me.serves.size() > B200_PLAN_MAX_HALO)
        return B200_ERR_UNSUPPORTED;
    const std::vector<Range> shares = split_rows(batch, B);
    *out = b200_rank_plan{};
    out->rank = rank; out->world = world; out->row_group = q; out->batch_group = bg;
    out->row_groups = R; out->batch_groups = B;
    out->img0 = shares[bg].lo; out->img1 = shares[bg].hi;
    out->out0 = me.out.lo; out->out1 = me.out.hi;
    out->conv2_0 = me.conv2.lo; out->conv2_1 = me.conv2.hi;
    out->need0 = me.need.lo; out->need1 = me.need.hi;
    out->own0 = me.own.lo; out->own1 = me.own.hi;
    out->conv1_0 = me.conv1.lo; out->conv1_1 = me.conv1.hi;
    out->in0 = me.in.lo; out->in1 = me.in.hi;
    out->n_halo = (int)me.halo.size(); out->n_serves = (int)me.serves.size();
    for (int i = 0; i < out->n_halo; ++i) {
        out->halo[i] = me.halo[i];
        out->halo[i].peer += bg * R;
    }
    for (int i = 0; i < out->n_serves; ++i) {
        out->serves[i] = me.serves[i];
        out->serves[i].peer += bg * R;
    }
    out->Hc1 = Hc1; out->Hp1 = Hp1; out->Hc2 = Hc2; out->Hp2 = Hp2;
    return B200_OK;
}
