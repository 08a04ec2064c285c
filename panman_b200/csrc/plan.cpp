// Plan builder: tree validation, chunking, walk-program emission.  Everything is iterative: ladders of depth 10^5 and
// more (BASELINE config 5) would overflow the call stack -- the reference's own recursion does.
#include "plan.hpp"

#include <algorithm>
#include <numeric>

namespace pmfs {
namespace {

inline uint32_t op(uint32_t code, uint32_t arg) { return code | (arg << kOpBits); }

}  // namespace

std::string build_plan(const pmfs_tree& t, const PlanOptions& opt, Plan& P) {
    int32_t chunk_nodes = opt.chunk_nodes;
    const int32_t PCH = std::min(std::max(opt.part_children, 1), kPartChildren);  // children per part of a split polytomy
    const int32_t V = t.n_nodes;
    if (V < 2) return "tree needs at least a root and one child";
    if (!t.parent || !t.child_off || !t.child_idx || !t.leaf_row) return "null tree array";
    if (t.parent[0] != -1) return "node 0 must be the root (parent -1)";
    if (t.child_off[0] != 0) return "child_off[0] must be 0";
    if ((int64_t)V >= (int64_t)1 << 28) return "too many nodes";
    for (int32_t v = 0; v < V; v++) {
        if (t.child_off[v + 1] < t.child_off[v]) return "child_off not monotone";
        if (v > 0 && (t.parent[v] < 0 || t.parent[v] >= v)) return "parent[i] must be in [0, i) (pre-order ids)";
    }
    if (t.child_off[V] != V - 1) return "child list must name every non-root node exactly once";
    if (t.child_off[1] == 0) return "root must be an internal node";
    {
        std::vector<uint8_t> seen(V, 0);
        for (int32_t v = 0; v < V; v++)
            for (int32_t e = t.child_off[v]; e < t.child_off[v + 1]; e++) {
                int32_t c = t.child_idx[e];
                if (c <= 0 || c >= V || seen[c] || t.parent[c] != v) return "child_idx inconsistent with parent[]";
                seen[c] = 1;
            }
    }

    P = Plan();
    P.n_nodes = V;
    P.int_id.assign(V, -1);
    int32_t max_row = -1;
    for (int32_t v = 0; v < V; v++) {
        int32_t k = t.child_off[v + 1] - t.child_off[v];
        P.max_arity = std::max(P.max_arity, k);
        if (k > 0) {
            if (t.leaf_row[v] != -1) return "leaf_row must be -1 for internal nodes";
            P.int_id[v] = P.n_internal++;
            P.int_to_node.push_back(v);
        } else {
            if (t.leaf_row[v] < -2 || t.leaf_row[v] == -1) return "leaf_row of a leaf must be a row >= 0 or -2";
            if ((int64_t)t.leaf_row[v] >= ((int64_t)1 << 28)) return "leaf_row too large";
            max_row = std::max(max_row, t.leaf_row[v]);
            P.n_leaves++;
        }
    }
    P.row_to_node.assign((size_t)max_row + 1, -1);
    for (int32_t v = 0; v < V; v++)
        if (t.leaf_row[v] >= 0) {
            // the reference keys every per-column map by NAME, so two leaves sharing a row would alias one entry
            if (P.row_to_node[t.leaf_row[v]] != -1) return "two leaves map to the same MSA row";
            P.row_to_node[t.leaf_row[v]] = v;
            P.n_rows_used++;
        }

    // ---- chunking: rem[v] = nodes of v's subtree not yet cut off; cut v once rem[v] >= threshold.  Round 0 cuts the
    // whole tree with `chunk_nodes`; every further round re-cuts only the residual top (nodes with a cut descendant)
    // with a threshold 8x smaller, until the residual is small.
    if (chunk_nodes < 2) chunk_nodes = 2;
    // A re-cut only pays where the residual top BRANCHES (the pieces below a branching node run in parallel); along a
    // path it would just chop a serial chain into more hand-overs -- a ladder tree stays in round-0 chunks.
    std::vector<int32_t> rem(V, 1);
    std::vector<uint8_t> cut(V, 0), hascut(V, 0), fork(V, 0);
    int64_t thr = chunk_nodes;
    for (int round = 0; round < 8; round++) {
        std::fill(rem.begin(), rem.end(), 1);
        if (round > 0) {  // fork[v]: at least two children of v lead to cuts of earlier rounds
            std::vector<uint8_t> n_hc(V, 0);
            for (int32_t v = 1; v < V; v++)
                if ((cut[v] || hascut[v]) && n_hc[t.parent[v]] < 2) n_hc[t.parent[v]]++;
            for (int32_t v = 0; v < V; v++) fork[v] = n_hc[v] >= 2;
        }
        bool any_new = false;
        for (int32_t v = V - 1; v >= 1; v--) {
            const bool candidate = P.int_id[v] >= 0 && !cut[v] && (round == 0 || (hascut[v] && fork[t.parent[v]]));
            if (candidate && rem[v] >= thr) {
                cut[v] = 1;
                any_new = true;
            }
            if (cut[v]) {
                rem[t.parent[v]] += 1;
                hascut[t.parent[v]] = 1;
            } else {
                rem[t.parent[v]] += rem[v];
                if (hascut[v]) hascut[t.parent[v]] = 1;
            }
        }
        if (!any_new || rem[0] < 32 || thr <= 8) break;
        thr = std::max<int64_t>(8, thr / 8);
    }
    // ---- split polytomies.  A node with more than split_arity children would be one serial stretch of one work item per
    // column tile (0.4 us per leaf child for Fitch, 2 us for Sankoff's counters): its children are dealt out, PCH
    // at a time in child-list order, to PART chunks that fold them into partial results side by side; the node's own
    // chunk only combines those.  Every internal child of such a node roots a chunk of its own, so that a part's
    // operands are leaves and finished sets only.
    std::vector<uint8_t> split(V, 0);
    if (opt.split_arity > 0)
        for (int32_t v = 0; v < V; v++)
            if (t.child_off[v + 1] - t.child_off[v] > opt.split_arity) {
                split[v] = 1;
                for (int32_t k = t.child_off[v]; k < t.child_off[v + 1]; k++)
                    if (P.int_id[t.child_idx[k]] >= 0) cut[t.child_idx[k]] = 1;
            }
    // ---- walk-stack depth.  Every stack level costs shared memory per CTA (16 KB bottom-up, 5 KB top-down) and with it
    // resident CTAs per SM, which is what the latency-bound walks live on; so a chunk may nest at most kMaxWalkStack
    // saved states.  Bottom-up a binary node with two in-chunk internal children saves the first one's set while it
    // walks the second; top-down every in-chunk internal child but the last saves the parent state.  Where that would
    // nest deeper, the child that would be walked with the stack already full is cut off as a chunk of its own.
    {
        constexpr int kMaxWalkStack = 2;
        std::vector<int32_t> remf(V, 1);
        std::vector<uint8_t> need_f(V, 0), need_b(V, 0);
        for (int32_t v = V - 1; v >= 0; v--) {
            if (P.int_id[v] < 0) continue;
            const int32_t b = t.child_off[v], e = t.child_off[v + 1];
            for (;;) {
                int32_t first = -1, last = -1, nsub = 0;  // the generators' orders: stable sorts on rem (see below)
                for (int32_t k = b; k < e; k++) {
                    const int32_t c = t.child_idx[k];
                    if (P.int_id[c] < 0 || cut[c]) continue;
                    nsub++;
                    if (first < 0 || remf[c] > remf[first]) first = c;
                    if (last < 0 || remf[c] >= remf[last]) last = c;
                }
                bool changed = false;
                int nf = 0, nb = 0;
                for (int32_t k = b; k < e; k++) {
                    const int32_t c = t.child_idx[k];
                    if (P.int_id[c] < 0 || cut[c]) continue;
                    const bool pushed_f = (e - b) == 2 && nsub == 2 && c != first;
                    const bool pushed_b = c != last;
                    if ((pushed_f && need_f[c] + 1 > kMaxWalkStack) || (pushed_b && need_b[c] + 1 > kMaxWalkStack)) {
                        cut[c] = 1;  // (all offenders of this scan at once: a node may have 10^5 children)
                        changed = true;
                        continue;
                    }
                    nf = std::max(nf, need_f[c] + (pushed_f ? 1 : 0));
                    nb = std::max(nb, need_b[c] + (pushed_b ? 1 : 0));
                }
                if (changed) continue;
                need_f[v] = (uint8_t)nf;
                need_b[v] = (uint8_t)nb;
                break;
            }
            for (int32_t k = b; k < e; k++) {
                const int32_t c = t.child_idx[k];
                remf[v] += (P.int_id[c] < 0 || cut[c]) ? 1 : remf[c];
            }
        }
    }
    // rem[] must describe the final cut set (the walk orders children by it)
    std::fill(rem.begin(), rem.end(), 1);
    for (int32_t v = V - 1; v >= 1; v--) rem[t.parent[v]] += cut[v] ? 1 : rem[v];
    cut[0] = 1;
    // chunk ids follow the pre-order node ids (a split node's own chunk, then its parts, then everything below): a chunk's
    // parent always has a smaller id
    std::vector<int32_t> chunk_of_root(V, -1), part_first(V, -1);
    for (int32_t v = 0; v < V; v++) {
        if (cut[v]) {
            chunk_of_root[v] = (int32_t)P.chunks.size();
            Chunk c;
            c.root_node = v;
            c.root_int = P.int_id[v];
            P.chunks.push_back(c);
        }
        if (split[v]) {
            part_first[v] = (int32_t)P.chunks.size();
            const int32_t k = t.child_off[v + 1] - t.child_off[v];
            for (int32_t p = 0; p * PCH < k; p++) {
                Chunk c;
                c.root_node = v;
                c.root_int = P.int_id[v];
                c.part_row = P.n_internal + (P.n_parts++) * opt.part_rows;
                P.chunks.push_back(c);
            }
        }
    }
    P.part_rows = P.n_parts ? opt.part_rows : 0;
    P.n_set_rows = P.n_internal + P.n_parts * P.part_rows;
    if ((int64_t)P.n_internal + (int64_t)P.n_parts * opt.part_rows >= (int64_t)kOperandIdMask) return "too many set rows";
    const int32_t NC = (int32_t)P.chunks.size();
    // owner[v]: chunk whose program visits v as a member (a chunk root is a member of its own chunk)
    std::vector<int32_t> owner(V, 0);
    for (int32_t v = 1; v < V; v++) owner[v] = cut[v] ? chunk_of_root[v] : owner[t.parent[v]];
    // the chunk that hands a chunk its parent state (and consumes its result): the owner of the root's parent -- or,
    // below a split node, the part that holds the root; a part hangs off the split node's own chunk
    std::vector<int32_t> child_pos;
    if (P.n_parts) {
        child_pos.assign(V, 0);
        for (int32_t v = 0; v < V; v++)
            if (split[v])
                for (int32_t k = t.child_off[v]; k < t.child_off[v + 1]; k++) child_pos[t.child_idx[k]] = k - t.child_off[v];
    }
    for (int32_t c = 1; c < NC; c++) {
        Chunk& ch = P.chunks[c];
        if (ch.part_row >= 0) {
            ch.parent = owner[ch.root_node];
        } else {
            const int32_t tp = t.parent[ch.root_node];
            ch.parent = split[tp] ? part_first[tp] + child_pos[ch.root_node] / PCH : owner[tp];
        }
        P.chunks[ch.parent].n_child++;
        ch.depth = P.chunks[ch.parent].depth + 1;
    }
    for (int32_t c = NC - 1; c >= 1; c--) {
        Chunk& par = P.chunks[P.chunks[c].parent];
        par.height = std::max(par.height, P.chunks[c].height + 1);
    }

    auto is_sub = [&](int32_t c) { return P.int_id[c] >= 0 && !cut[c]; };
    auto operand = [&](int32_t c) -> uint32_t {
        if (P.int_id[c] >= 0) return kOperandSet | (cut[c] ? kOperandExt : 0u) | (uint32_t)P.int_id[c];
        return t.leaf_row[c] >= 0 ? (uint32_t)t.leaf_row[c] : kAbsentLeaf;
    };
    // pushes an operand word on the program and the loads it will consume on the stream
    auto push_operand = [&](int32_t c) {
        const uint32_t o = operand(c);
        P.fwd_prog.push_back(o);
        if (o & kOperandSet) {
            if (o & kOperandExt) P.fwd_stream.push_back(kStreamSet | (o & kOperandIdMask));
        } else if (o != kAbsentLeaf) {
            P.fwd_stream.push_back(o);
        }
    };

    // ---- forward (post-order) programs
    struct Frame {
        int32_t v;
        int32_t i;
        bool pushed;
        std::vector<int32_t> subs;
    };
    std::vector<Frame> st;
    for (int32_t ci = 0; ci < NC; ci++) {
        Chunk& ch = P.chunks[ci];
        while (P.fwd_prog.size() % 4) P.fwd_prog.push_back(0);  // 16-byte aligned: programs are staged by bulk copies
        ch.fwd_off = (int64_t)P.fwd_prog.size();
        ch.fwd_stream_off = (int64_t)P.fwd_stream.size();
        int32_t depth = 0;
        st.clear();
        if (ch.part_row >= 0) {  // a part: one fold over its share of the split node's children
            const int32_t v = ch.root_node, p = ci - part_first[v];
            const int32_t b = t.child_off[v] + p * PCH, e = std::min(t.child_off[v + 1], b + PCH);
            P.fwd_prog.push_back(op(F_PART, (uint32_t)ch.part_row));
            P.fwd_prog.push_back((uint32_t)(e - b));
            for (int32_t k = b; k < e; k++) push_operand(t.child_idx[k]);
        } else {
            st.push_back({ch.root_node, -1, false, {}});
        }
        while (!st.empty()) {
            Frame& f = st.back();
            const int32_t v = f.v, b = t.child_off[v], e = t.child_off[v + 1];
            if (split[v]) {  // the parts have folded the children: combine their partial results
                const int32_t m = (e - b + PCH - 1) / PCH;
                P.fwd_prog.push_back(op(F_COMBINE, (uint32_t)P.int_id[v]));
                P.fwd_prog.push_back((uint32_t)m);
                for (int32_t p = 0; p < m; p++) {
                    const uint32_t row = (uint32_t)P.chunks[part_first[v] + p].part_row;
                    P.fwd_prog.push_back(kOperandSet | kOperandExt | row);
                    for (int32_t r = 0; r < P.part_rows; r++) P.fwd_stream.push_back(kStreamSet | (row + (uint32_t)r));
                }
                st.pop_back();
                continue;
            }
            if (f.i < 0) {
                for (int32_t k = b; k < e; k++)
                    if (is_sub(t.child_idx[k])) f.subs.push_back(t.child_idx[k]);
                std::stable_sort(f.subs.begin(), f.subs.end(), [&](int32_t x, int32_t y) { return rem[x] > rem[y]; });
                f.i = 0;
            }
            const bool binary = (e - b) == 2;
            const int32_t ns = (int32_t)f.subs.size();
            if (binary && ns == 2) {
                if (f.i == 0) {
                    f.i = 1;
                    int32_t c = f.subs[0];
                    st.push_back({c, -1, false, {}});
                } else if (f.i == 1) {
                    P.fwd_prog.push_back(op(F_PUSH, 0));
                    P.fwd_stack_depth = std::max(P.fwd_stack_depth, ++depth);
                    f.i = 2;
                    int32_t c = f.subs[1];
                    st.push_back({c, -1, false, {}});
                } else {
                    P.fwd_prog.push_back(op(F_JOIN_POP, (uint32_t)P.int_id[v]));
                    depth--;
                    st.pop_back();
                }
            } else if (binary && ns == 1) {
                if (f.i == 0) {
                    f.i = 1;
                    int32_t c = f.subs[0];
                    st.push_back({c, -1, false, {}});
                } else {
                    int32_t other = t.child_idx[b] == f.subs[0] ? t.child_idx[b + 1] : t.child_idx[b];
                    P.fwd_prog.push_back(op(F_JOIN_IMM, (uint32_t)P.int_id[v]));
                    push_operand(other);
                    st.pop_back();
                }
            } else if (binary) {
                P.fwd_prog.push_back(op(F_JOIN_IMM2, (uint32_t)P.int_id[v]));
                push_operand(t.child_idx[b]);
                push_operand(t.child_idx[b + 1]);
                st.pop_back();
            } else {
                if (f.i < ns) {
                    int32_t c = f.subs[f.i++];
                    st.push_back({c, -1, false, {}});
                } else {
                    P.fwd_prog.push_back(op(F_POLY, (uint32_t)P.int_id[v]));
                    P.fwd_prog.push_back((uint32_t)(e - b));
                    for (int32_t k = b; k < e; k++) push_operand(t.child_idx[k]);
                    st.pop_back();
                }
            }
        }
        ch.fwd_len = (int32_t)((int64_t)P.fwd_prog.size() - ch.fwd_off);
        ch.fwd_stream_len = (int32_t)((int64_t)P.fwd_stream.size() - ch.fwd_stream_off);
    }

    // ---- backward (pre-order) programs
    for (int32_t ci = 0; ci < NC; ci++) {
        Chunk& ch = P.chunks[ci];
        while (P.bwd_prog.size() % 4) P.bwd_prog.push_back(0);
        ch.bwd_off = (int64_t)P.bwd_prog.size();
        ch.bwd_stream_off = (int64_t)P.bwd_stream.size();
        int32_t depth = 0;
        st.clear();
        if (ch.part_row >= 0) {  // a part starts from the split node's state: its share of the children, child chunks first
            const int32_t v = ch.root_node, p = ci - part_first[v];
            const int32_t b = t.child_off[v] + p * PCH, e = std::min(t.child_off[v + 1], b + PCH);
            for (int32_t k = b; k < e; k++)
                if (P.int_id[t.child_idx[k]] >= 0) P.bwd_prog.push_back(op(B_XCHILD, (uint32_t)chunk_of_root[t.child_idx[k]]));
            for (int32_t k = b; k < e; k++) {
                const int32_t c = t.child_idx[k];
                if (P.int_id[c] < 0 && t.leaf_row[c] >= 0) {
                    P.bwd_prog.push_back(op(B_LEAF, (uint32_t)t.leaf_row[c]));
                    P.bwd_stream.push_back((uint32_t)t.leaf_row[c]);
                }
            }
            if (ch.bwd_off == (int64_t)P.bwd_prog.size()) {  // (absent leaves only: nothing to do, but no item is empty)
                P.bwd_prog.push_back(op(B_PUSHP, 0));
                P.bwd_prog.push_back(op(B_POPP, 0));
                P.bwd_stack_depth = std::max(P.bwd_stack_depth, 1);
            }
        } else {
            st.push_back({ch.root_node, -1, false, {}});
        }
        while (!st.empty()) {
            Frame& f = st.back();
            const int32_t v = f.v, b = t.child_off[v], e = t.child_off[v + 1];
            if (split[v]) {  // resolve the node, hand its state to the parts
                P.bwd_prog.push_back(v == 0 ? op(B_ENTER_ROOT, 0) : op(B_ENTER, (uint32_t)P.int_id[v]));
                P.bwd_stream.push_back(kStreamSet | (uint32_t)P.int_id[v]);
                for (int32_t p = 0; p * PCH < e - b; p++) P.bwd_prog.push_back(op(B_XCHILD, (uint32_t)(part_first[v] + p)));
                st.pop_back();
                continue;
            }
            if (f.i < 0) {
                P.bwd_prog.push_back(v == 0 ? op(B_ENTER_ROOT, 0) : op(B_ENTER, (uint32_t)P.int_id[v]));
                P.bwd_stream.push_back(kStreamSet | (uint32_t)P.int_id[v]);
                // child chunks first: XCHILD releases the child chunk (its tile can start at once), so it comes before
                // anything that takes time at this node
                for (int32_t k = b; k < e; k++) {
                    int32_t c = t.child_idx[k];
                    if (P.int_id[c] >= 0 && cut[c]) P.bwd_prog.push_back(op(B_XCHILD, (uint32_t)chunk_of_root[c]));
                }
                for (int32_t k = b; k < e; k++) {
                    int32_t c = t.child_idx[k];
                    if (P.int_id[c] < 0) {
                        if (t.leaf_row[c] >= 0) {
                            P.bwd_prog.push_back(op(B_LEAF, (uint32_t)t.leaf_row[c]));
                            P.bwd_stream.push_back((uint32_t)t.leaf_row[c]);
                        }
                    } else if (!cut[c]) {
                        f.subs.push_back(c);
                    }
                }
                std::stable_sort(f.subs.begin(), f.subs.end(), [&](int32_t x, int32_t y) { return rem[x] < rem[y]; });
                f.i = 0;
            }
            if (f.pushed) {
                P.bwd_prog.push_back(op(B_POPP, 0));
                depth--;
                f.pushed = false;
            }
            if (f.i < (int32_t)f.subs.size()) {
                int32_t c = f.subs[f.i];
                bool last = f.i + 1 == (int32_t)f.subs.size();
                f.i++;
                if (!last) {
                    P.bwd_prog.push_back(op(B_PUSHP, 0));
                    P.bwd_stack_depth = std::max(P.bwd_stack_depth, ++depth);
                    f.pushed = true;
                }
                st.push_back({c, -1, false, {}});  // invalidates f
            } else {
                st.pop_back();
            }
        }
        ch.bwd_len = (int32_t)((int64_t)P.bwd_prog.size() - ch.bwd_off);
        ch.bwd_stream_len = (int32_t)((int64_t)P.bwd_stream.size() - ch.bwd_stream_off);
    }

    for (int i = 0; i < 4; i++) {  // bulk copies round a program's length up to 16 bytes
        P.fwd_prog.push_back(0);
        P.bwd_prog.push_back(0);
    }

    // ---- execution order: critical-path list scheduling.  Any order in which dependencies come first is valid (the
    // kernels claim items in this order and wait for unfinished dependencies).  Bottom-up: a chunk's longest remaining
    // chain is its distance to the root chunk, so DEEPEST chunks go first -- the long chains get going at once and the
    // many shallow leaf chunks are the filler that keeps the SMs busy while the chains climb.  Top-down: the longest
    // remaining chain is the chunk's height, so TALLEST first.  (Plain level order leaves the whole chain of narrow
    // levels near the root exposed at the end / start of a pass.)  lvl_off only counts chunks per height / depth.
    // remaining chain length in program words: up[c] = own + up[parent] (bottom-up pass), down[c] = own + max down[child]
    std::vector<int64_t> up(NC, 0), down(NC, 0);
    for (int32_t c = 0; c < NC; c++) up[c] = P.chunks[c].fwd_len + (P.chunks[c].parent >= 0 ? up[P.chunks[c].parent] : 0);
    for (int32_t c = NC - 1; c >= 0; c--) {
        down[c] += P.chunks[c].bwd_len;
        if (P.chunks[c].parent >= 0) down[P.chunks[c].parent] = std::max(down[P.chunks[c].parent], down[c]);
    }
    auto make_levels = [&](bool fwd, std::vector<int32_t>& order, std::vector<int32_t>& off) {
        int32_t nl = 0;
        for (auto& c : P.chunks) nl = std::max(nl, (fwd ? c.height : c.depth) + 1);
        order.resize(NC);
        std::iota(order.begin(), order.end(), 0);
        std::stable_sort(order.begin(), order.end(), [&](int32_t x, int32_t y) {
            const int64_t px = fwd ? up[x] : down[x], py = fwd ? up[y] : down[y];
            if (px != py) return px > py;
            int32_t wx = fwd ? P.chunks[x].fwd_len : P.chunks[x].bwd_len, wy = fwd ? P.chunks[y].fwd_len : P.chunks[y].bwd_len;
            return wx > wy;
        });
        off.assign(nl + 1, 0);
        for (auto& c : P.chunks) off[(fwd ? c.height : c.depth) + 1]++;
        for (int32_t l = 0; l < nl; l++) off[l + 1] += off[l];
    };
    make_levels(true, P.fwd_order, P.fwd_lvl_off);
    make_levels(false, P.bwd_order, P.bwd_lvl_off);

    for (int32_t v = 1; v < V; v++) (P.int_id[v] >= 0 ? P.edges_internal : P.edges_leaf)++;
    return "";
}

}  // namespace pmfs
