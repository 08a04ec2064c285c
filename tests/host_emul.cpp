// TEST INFRASTRUCTURE ONLY.  Runs the library's own plan builder (panman_b200/csrc/plan.cpp) and walk interpreter
// (panman_b200/csrc/walk.cuh, the __host__ side of the very code the CUDA kernels execute per thread) on the CPU, one
// column group at a time, so that program emission and the bit-sliced set algebra can be checked against the oracle
// without a GPU.  It is built by tests/test_host_emul.py into tests/_build/ and is never loaded by the product.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../panman_b200/csrc/plan.hpp"
#include "../panman_b200/csrc/walk.cuh"

using namespace pmfs;

namespace {

struct Rec {
    int32_t node, pos;
    uint8_t type, code, parent_code, pad;
};

template <int NG>
struct HostEnv {
    static constexpr int kGroups = NG;
    const Plan* plan;
    const uint8_t* msa;
    int64_t row_stride;
    int64_t G;
    int64_t gs[NG];  // clamped column groups this "thread" walks
    bool act[NG];    // false: padding lane (loads clamped, nothing stored or emitted)
    int64_t col_base;
    std::vector<uint32_t>*S, *D;       // [I][16][G] words
    std::vector<uint32_t>*xcode, *xv;  // [NC][4][G], [NC][G]
    std::vector<std::vector<Planes16>> fstk;
    std::vector<std::pair<std::vector<Code4>, std::vector<uint32_t>>> bstk;
    std::vector<Rec>* out;
    const uint32_t* stream = nullptr;  // the chunk's load stream: every ring-served load must match it, in order
    int32_t stream_len = 0, consumed = 0;
    bool stream_ok = true;

    void expect(uint32_t entry) {
        if (consumed >= stream_len || stream[consumed] != entry) stream_ok = false;
        consumed++;
    }
    uint32_t word(const uint32_t* prog, int32_t pc) const { return prog[pc]; }
    void leaf(uint32_t row, uint4 (&w)[NG]) {
        expect(row);
        for (int j = 0; j < NG; j++) memcpy(&w[j], msa + (int64_t)row * row_stride + gs[j] * 16, 16);
    }
    void load(std::vector<uint32_t>* a, uint32_t id, Planes16 (&s)[NG]) const {
        for (int j = 0; j < NG; j++)
            for (int i = 0; i < 16; i++) s[j].p[i] = (*a)[((int64_t)id * 16 + i) * G + gs[j]];
    }
    void store(std::vector<uint32_t>* a, uint32_t id, const Planes16 (&s)[NG]) const {
        for (int j = 0; j < NG; j++)
            if (act[j])
                for (int i = 0; i < 16; i++) (*a)[((int64_t)id * 16 + i) * G + gs[j]] = s[j].p[i];
    }
    void load_x(uint32_t id, Planes16 (&s)[NG]) {
        expect(kStreamSet | id);
        load(S, id, s);
    }
    void load_s(uint32_t id, Planes16 (&s)[NG]) const { load(S, id, s); }
    void load_d(uint32_t id, Planes16 (&s)[NG]) const { load(D, id, s); }
    void store_s(uint32_t id, const Planes16 (&s)[NG]) const { store(S, id, s); }
    void store_d(uint32_t id, const Planes16 (&s)[NG]) const { store(D, id, s); }
    void push(const Planes16 (&s)[NG]) { fstk.emplace_back(s, s + NG); }
    void pop(Planes16 (&s)[NG]) {
        for (int j = 0; j < NG; j++) s[j] = fstk.back()[j];
        fstk.pop_back();
    }
    void pushp(const Code4 (&c)[NG], const uint32_t (&pv)[NG]) {
        bstk.push_back({std::vector<Code4>(c, c + NG), std::vector<uint32_t>(pv, pv + NG)});
    }
    void popp(Code4 (&c)[NG], uint32_t (&pv)[NG]) {
        for (int j = 0; j < NG; j++) {
            c[j] = bstk.back().first[j];
            pv[j] = bstk.back().second[j];
        }
        bstk.pop_back();
    }
    void store_x(uint32_t chunk, const Code4 (&c)[NG], const uint32_t (&pv)[NG]) const {
        for (int j = 0; j < NG; j++) {
            if (!act[j]) continue;
            for (int i = 0; i < 4; i++) (*xcode)[((int64_t)chunk * 4 + i) * G + gs[j]] = c[j].b[i];
            (*xv)[(int64_t)chunk * G + gs[j]] = pv[j];
        }
    }
    bool any(bool x) const { return x; }
    void emit(int32_t node, const uint32_t (&m)[NG], const Code4 (&Gc)[NG], const Code4 (&P)[NG]) const {
        for (int j = 0; j < NG; j++)
            for (int q = 0; q < 32; q++)
                if ((m[j] >> q) & 1) {
                    uint32_t cg = 0, cp = 0;
                    for (int i = 0; i < 4; i++) {
                        cg |= ((Gc[j].b[i] >> q) & 1) << i;
                        cp |= ((P[j].b[i] >> q) & 1) << i;
                    }
                    Rec r;
                    r.node = node;
                    r.pos = (int32_t)(col_base + gs[j] * 32 + column_of_bit(q));
                    r.type = cp == 0 ? PMFS_NI : (cg == 0 ? PMFS_ND : PMFS_NS);
                    r.code = (uint8_t)cg;
                    r.parent_code = (uint8_t)cp;
                    r.pad = 0;
                    out->push_back(r);
                }
    }
    void emit_internal(uint32_t id, const uint32_t (&m)[NG], const Code4 (&Gc)[NG], const Code4 (&P)[NG]) const {
        emit(plan->int_to_node[id], m, Gc, P);
    }
    void emit_leaf(uint32_t row, const uint32_t (&m)[NG], const Code4 (&Gc)[NG], const Code4 (&P)[NG]) const {
        emit(plan->row_to_node[row], m, Gc, P);
    }
};

Code4 planes_of(const uint8_t* codes, int64_t n_cols, int64_t g) {
    Code4 c = {{0, 0, 0, 0}};
    for (int q = 0; q < 32; q++) {
        int64_t col = g * 32 + column_of_bit(q);
        uint32_t x = col < n_cols ? (codes[col] & 15u) : 0u;
        for (int i = 0; i < 4; i++) c.b[i] |= ((x >> i) & 1u) << q;
    }
    return c;
}

template <int ALGO, int NG>
int run(const Plan& P, const uint8_t* msa, int64_t row_stride, int64_t n_cols, const uint8_t* cons, const uint8_t* forced,
        std::vector<Rec>& out) {
    const int64_t G = (n_cols + 31) / 32;
    const int64_t H = (G + NG - 1) / NG;  // "threads": thread t walks groups t, t+H, ...
    std::vector<uint32_t> S((size_t)P.n_set_rows * 16 * G), D(ALGO ? (size_t)P.n_internal * 16 * G : 0);
    std::vector<uint32_t> xcode(P.chunks.size() * 4 * G), xv(P.chunks.size() * G);
    uint32_t err = 0;
    for (int pass = 0; pass < 2; pass++) {
        const auto& lvl = pass == 0 ? P.fwd_lvl_off : P.bwd_lvl_off;
        const auto& order = pass == 0 ? P.fwd_order : P.bwd_order;
        for (size_t l = 0; l + 1 < lvl.size(); l++)
            for (int32_t k = lvl[l]; k < lvl[l + 1]; k++) {
                const Chunk& ch = P.chunks[order[k]];
                for (int64_t t = 0; t < H; t++) {
                    HostEnv<NG> env{&P, msa, row_stride, G, {}, {}, 0, &S, &D, &xcode, &xv, {}, {}, &out};
                    for (int j = 0; j < NG; j++) {
                        const int64_t g = t + j * H;
                        env.act[j] = g < G;
                        env.gs[j] = g < G ? g : G - 1;
                    }
                    env.stream = pass == 0 ? P.fwd_stream.data() + ch.fwd_stream_off : P.bwd_stream.data() + ch.bwd_stream_off;
                    env.stream_len = pass == 0 ? ch.fwd_stream_len : ch.bwd_stream_len;
                    Code4 fc[NG];
                    const Code4* fp = nullptr;
                    if (forced) {
                        for (int j = 0; j < NG; j++) fc[j] = planes_of(forced, n_cols, env.gs[j]);
                        fp = fc;
                    }
                    if (pass == 0) {
                        forward_walk<ALGO, true>(env, P.fwd_prog.data() + ch.fwd_off, ch.fwd_len, ALGO == kAlgoFitch ? fp : nullptr);
                        if (!env.fstk.empty()) return -100;
                        if (!env.stream_ok || env.consumed != env.stream_len) return -102;
                    } else {
                        Code4 Pc[NG];
                        uint32_t pv[NG], colmask[NG];
                        for (int j = 0; j < NG; j++) {
                            const int64_t g = env.gs[j];
                            pv[j] = ~0u;
                            if (order[k] == 0) {
                                Pc[j] = planes_of(cons, n_cols, g);
                            } else {
                                for (int i = 0; i < 4; i++) Pc[j].b[i] = xcode[((int64_t)order[k] * 4 + i) * G + g];
                                pv[j] = xv[(int64_t)order[k] * G + g];
                            }
                            colmask[j] = 0;
                            if (env.act[j])
                                for (int q = 0; q < 32; q++)
                                    if (g * 32 + column_of_bit(q) < n_cols) colmask[j] |= 1u << q;
                        }
                        err |= backward_walk<ALGO>(env, P.bwd_prog.data() + ch.bwd_off, ch.bwd_len, Pc, pv,
                                                   ALGO == kAlgoSankoff ? fp : nullptr, colmask);
                        if (!env.bstk.empty()) return -101;
                        if (!env.stream_ok || env.consumed != env.stream_len) return -103;
                    }
                }
            }
    }
    if (err) return PMFS_ERR_ROOT_INVALID;
    std::sort(out.begin(), out.end(), [](const Rec& a, const Rec& b) { return a.node != b.node ? a.node < b.node : a.pos < b.pos; });
    return 0;
}

}  // namespace

static int32_t g_split_arity = 128, g_part_children = kPartChildren;

extern "C" {

void emul_free(void* p) { free(p); }

// plan options for split polytomies (PlanOptions::split_arity / part_children) of the following emul_run calls
void emul_set_split(int32_t split_arity, int32_t part_children) {
    g_split_arity = split_arity;
    g_part_children = part_children;
}

// same contract as pmfs_run (include/pmfs.h) + an explicit chunk size and the number of column groups per thread; plan_out (optional) receives
// {n_chunks, fwd_levels, bwd_levels, fwd_stack, bwd_stack, program_words}
int emul_run(int algo, int32_t n_nodes, const int32_t* parent, const int32_t* child_off, const int32_t* child_idx,
             const int32_t* leaf_row, const uint8_t* msa, int64_t row_stride, int64_t n_cols, const uint8_t* cons,
             const uint8_t* forced, int32_t chunk_nodes, int32_t groups_per_thread, void** out, int64_t* n_out,
             int64_t* plan_out) {
    pmfs_tree t = {n_nodes, parent, child_off, child_idx, leaf_row};
    Plan P;
    PlanOptions po;
    po.chunk_nodes = chunk_nodes;
    po.split_arity = g_split_arity;
    po.part_children = g_part_children;
    po.part_rows = algo == 0 ? kPartRowsFitch : kPartRowsSankoff;
    std::string e = build_plan(t, po, P);
    if (!e.empty()) return PMFS_ERR_ARG;
    if (plan_out) {
        plan_out[0] = (int64_t)P.chunks.size();
        plan_out[1] = (int64_t)P.fwd_lvl_off.size() - 1;
        plan_out[2] = (int64_t)P.bwd_lvl_off.size() - 1;
        plan_out[3] = P.fwd_stack_depth;
        plan_out[4] = P.bwd_stack_depth;
        plan_out[5] = (int64_t)(P.fwd_prog.size() + P.bwd_prog.size());
        plan_out[6] = P.n_parts;
    }
    // rows padded to whole groups, as the library does on upload
    const int64_t dstride = ((n_cols + 31) / 32) * 16;
    int32_t n_rows = (int32_t)P.row_to_node.size();
    std::vector<uint8_t> padded((size_t)std::max(1, n_rows) * dstride, 0);
    for (int32_t r = 0; r < n_rows; r++) memcpy(padded.data() + (size_t)r * dstride, msa + (size_t)r * row_stride, (size_t)((n_cols + 1) / 2));
    std::vector<Rec> recs;
    int rc;
    if (groups_per_thread == 2)
        rc = algo == 0 ? run<kAlgoFitch, 2>(P, padded.data(), dstride, n_cols, cons, forced, recs)
                       : run<kAlgoSankoff, 2>(P, padded.data(), dstride, n_cols, cons, forced, recs);
    else
        rc = algo == 0 ? run<kAlgoFitch, 1>(P, padded.data(), dstride, n_cols, cons, forced, recs)
                       : run<kAlgoSankoff, 1>(P, padded.data(), dstride, n_cols, cons, forced, recs);
    if (rc) return rc;
    Rec* o = (Rec*)malloc(std::max<size_t>(1, recs.size()) * sizeof(Rec));
    if (!recs.empty()) memcpy(o, recs.data(), recs.size() * sizeof(Rec));
    *out = o;
    *n_out = (int64_t)recs.size();
    return 0;
}
}
