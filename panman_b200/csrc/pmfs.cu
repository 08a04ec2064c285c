// libpmfs: context, pipeline and C ABI (include/pmfs.h) around the CUDA kernels (sm_100a).
//
// Device code:
//   walk_kernels.cuh  fwd_kernel<CFG,ALGO,POLY>  bottom-up walk, ONE launch: persistent CTAs claim (chunk, column tile) items
//                     bwd_kernel<CFG,ALGO>       top-down walk + mutation emission (warp-cooperative stream compaction)
//                     templates on the tile shape CFG, instantiated once per shape in walk_inst_*.cu and reached from
//                     here through walk_params.hpp's WalkKernelSet; they run walk.cuh's interpreter (shared with the
//                     host emulation in tests/)
//   aux_kernels.cuh   (included below) prep_planes_kernel (codes -> bit-sliced planes), order_cells_kernel + helpers
//                     (ordering without a sort), compact/assemble (radix path), pack_* (<=6-run packing),
//                     pack_chars_kernel, synth_* kernels
// HBM layout: MSA rows of packed nibbles (16 B per thread per leaf); set arrays S (and D1 for Sankoff)
// tile-major, i.e. one node's set for one column tile is one contiguous block of up to 16 KB (a single TMA copy).
#include <cuda_runtime.h>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/pmfs.h"
#include "plan.hpp"
#include "walk.cuh"
#include "walk_params.hpp"
#include "ctx.hpp"

namespace pmfs {

// A mutation record is ONE 64-bit word: sort key (node, column) in the high bits, payload in the low kPayloadBits = 10
// bits (type 2 bits, node state 4, parent state 4; walk_params.hpp), so ordering the records is a keys-only radix sort
// over exactly the key bits that can be set, and the multi-GPU exchange moves 8 bytes per record.

thread_local std::string g_error;

}  // namespace pmfs

#include "aux_kernels.cuh"

// ---- tile shapes (walk_inst_*.cu: one translation unit per shape) ------------------------------------------------------
namespace pmfs {
const WalkKernelSet* walk_kernel_set_4x2();
const WalkKernelSet* walk_kernel_set_2x2();
const WalkKernelSet* walk_kernel_set_1x2();
const WalkKernelSet* walk_kernel_set_4x1();
const WalkKernelSet* walk_kernel_set_2x1();
const WalkKernelSet* walk_kernel_set_1x1();
const WalkKernelSet* walk_kernel_set_by_index(int i) {
    switch (i) {
        case 0: return walk_kernel_set_4x2();  // 256 column groups per work item
        case 1: return walk_kernel_set_2x2();  // 128
        case 2: return walk_kernel_set_4x1();  // 128, one group per thread
        case 3: return walk_kernel_set_1x2();  // 64
        case 4: return walk_kernel_set_2x1();  // 64, one group per thread
        case 5: return walk_kernel_set_1x1();  // 32
        default: return nullptr;
    }
}
const WalkKernelSet* walk_kernel_set(int nw, int ng) {
    for (int i = 0; i < kNumTileShapes; i++) {
        const WalkKernelSet* s = walk_kernel_set_by_index(i);
        if (s->shape.nw == nw && s->shape.ng == ng) return s;
    }
    return nullptr;
}
}  // namespace pmfs

static int upload_plan(pmfs_ctx* c) {
    const Plan& P = c->plan;
    const size_t NC = P.chunks.size();
    std::vector<int64_t> foff(NC), boff(NC), fsoff(NC), bsoff(NC);
    std::vector<int32_t> flen(NC), blen(NC), fslen(NC), bslen(NC), cpar(NC), cnch(NC);
    for (size_t i = 0; i < NC; i++) {
        foff[i] = P.chunks[i].fwd_off; flen[i] = P.chunks[i].fwd_len;
        boff[i] = P.chunks[i].bwd_off; blen[i] = P.chunks[i].bwd_len;
        fsoff[i] = P.chunks[i].fwd_stream_off; fslen[i] = P.chunks[i].fwd_stream_len;
        bsoff[i] = P.chunks[i].bwd_stream_off; bslen[i] = P.chunks[i].bwd_stream_len;
        cpar[i] = P.chunks[i].parent; cnch[i] = P.chunks[i].n_child;
    }
    auto up = [&](auto& buf, const auto& v) -> int {
        int rc = buf.ensure(v.size());
        if (rc) return rc;
        if (!v.empty()) PMFS_CUDA(cudaMemcpy(buf.ptr, v.data(), v.size() * sizeof(v[0]), cudaMemcpyHostToDevice));
        return PMFS_OK;
    };
    int rc;
    if ((rc = up(c->d_fwd_prog, P.fwd_prog))) return rc;
    if ((rc = up(c->d_bwd_prog, P.bwd_prog))) return rc;
    for (int pass = 0; pass < 2; pass++) {  // packed per-item records in execution order (see WalkParams::items)
        const std::vector<int32_t>& order = pass == 0 ? P.fwd_order : P.bwd_order;
        std::vector<int4> items(2 * NC);
        for (size_t k = 0; k < NC; k++) {
            const Chunk& ch = P.chunks[order[k]];
            const int64_t poff = pass == 0 ? ch.fwd_off : ch.bwd_off, soff = pass == 0 ? ch.fwd_stream_off : ch.bwd_stream_off;
            if (poff / 4 > 0x7fffffffll || soff > 0x7fffffffll) {
                g_error = "walk programs exceed 2^31 words";
                return PMFS_ERR_ARG;
            }
            items[2 * k] = make_int4(order[k], ch.parent, pass == 0 ? ch.fwd_len : ch.bwd_len,
                                     pass == 0 ? ch.fwd_stream_len : ch.bwd_stream_len);
            // both passes wait on the chunk's OWN counter: bottom-up it counts finished child chunks, top-down it is set
            // by the parent chunk the moment that has stored this chunk's parent state (XCHILD), not when it finishes
            const int need = pass == 0 ? ch.n_child : (ch.parent < 0 ? 0 : 1);
            const int dep = order[k];
            items[2 * k + 1] = make_int4((int)(poff / 4), (int)soff, need, dep);
        }
        if ((rc = up(pass == 0 ? c->d_fwd_items : c->d_bwd_items, items))) return rc;
    }
    if ((rc = up(c->d_chunk_parent, cpar))) return rc;
    if ((rc = up(c->d_chunk_nchild, cnch))) return rc;
    if ((rc = up(c->d_fwd_stream, P.fwd_stream))) return rc;
    if ((rc = up(c->d_bwd_stream, P.bwd_stream))) return rc;
    if ((rc = up(c->d_fwd_soff, fsoff))) return rc;
    if ((rc = up(c->d_bwd_soff, bsoff))) return rc;
    if ((rc = up(c->d_fwd_slen, fslen))) return rc;
    if ((rc = up(c->d_bwd_slen, bslen))) return rc;
    if ((rc = up(c->d_fwd_off, foff))) return rc;
    if ((rc = up(c->d_bwd_off, boff))) return rc;
    if ((rc = up(c->d_fwd_len, flen))) return rc;
    if ((rc = up(c->d_bwd_len, blen))) return rc;
    if ((rc = up(c->d_fwd_order, P.fwd_order))) return rc;
    if ((rc = up(c->d_bwd_order, P.bwd_order))) return rc;
    if ((rc = up(c->d_int_to_node, P.int_to_node))) return rc;
    if ((rc = up(c->d_row_to_node, P.row_to_node))) return rc;
    return PMFS_OK;
}

static int replan(pmfs_ctx* c, int32_t chunk_nodes, int algo = -1) {
    pmfs_tree t;
    t.n_nodes = (int32_t)c->parent.size();
    t.parent = c->parent.data();
    t.child_off = c->child_off.data();
    t.child_idx = c->child_idx.data();
    t.leaf_row = c->leaf_row.data();
    // (the programs of split polytomies depend on the algorithm: a part's partial result is 2 set rows for Fitch, 6 for
    // Sankoff; without an algorithm given, the one the current plan was built for)
    PlanOptions po;
    po.chunk_nodes = chunk_nodes;
    po.split_arity = c->tune_split_arity;
    po.part_children = c->tune_part_children;
    po.part_rows = algo == PMFS_SANKOFF || (algo < 0 && c->plan.part_rows == kPartRowsSankoff) ? kPartRowsSankoff : kPartRowsFitch;
    std::string err = build_plan(t, po, c->plan);
    if (!err.empty()) {
        g_error = "pmfs_create: " + err;
        return PMFS_ERR_ARG;
    }
    c->has_poly = false;
    for (int32_t v = 0; v < t.n_nodes; v++)
        if (t.child_off[v + 1] - t.child_off[v] != 0 && t.child_off[v + 1] - t.child_off[v] != 2) c->has_poly = true;
    c->chunk_nodes_used = chunk_nodes;
    return upload_plan(c);
}

// Chunk size.  At BASELINE-config-2 scale a pass is (chain of dependent items near the root) + (bulk throughput).  With
// the walk stack capped by the plan (plan.cpp: kMaxWalkStack, so that chunk size no longer costs resident CTAs) and child
// chunks released per XCHILD, the fixed cost per item is what is left to amortise: measured on config 2 the step is flat
// between 24 and 64 nodes (0.71-0.72 ms) and 4 % slower at 14 (sweeps in profiles/README.md).
static int32_t auto_chunk_nodes(int32_t n_nodes, int64_t groups, int tile_groups) {
    int64_t ctas_per_chunk = (groups + tile_groups - 1) / tile_groups;
    int64_t want_chunks = std::max<int64_t>(1, (148 * 48 + ctas_per_chunk - 1) / ctas_per_chunk);
    int64_t cn = n_nodes / (want_chunks * 2);  // chunks end up between cn and 2*cn nodes
    return (int32_t)std::min<int64_t>(std::max<int64_t>(cn, 24), 1 << 16);
}

// Tile shape of a run over `groups` column groups (per memory tile).  Wide tiles amortise per-item and per-op
// bookkeeping and are what BASELINE config 2 was tuned on; narrow ones are for (1) trees that offer no chunk-level
// parallelism -- on a ladder the column tiles are the only independent work, so the narrowest tile gives the most
// chains --, and (2) narrow column ranges (column shards of a wide tree, memory-limited column batches), where a wide
// tile would idle most of its lanes.  PMFS_TILE_SHAPE=NWxNG (environment) and pmfs_set_tile override the choice.
// max_tile_groups > 0: the set arrays only fit that many column groups per pass, and a pass is a whole number of tiles --
// shapes wider than that are out (rounding the pass up to one tile would allocate more than what fits).
static const WalkKernelSet* choose_shape(const pmfs_ctx* c, int64_t groups, int64_t max_tile_groups = 0) {
    if (c->tune_nw > 0)
        if (const WalkKernelSet* s = walk_kernel_set(c->tune_nw, c->tune_ng)) return s;
    if (const char* e = getenv("PMFS_TILE_SHAPE")) {
        int nw = 0, ng = 0;
        if (sscanf(e, "%dx%d", &nw, &ng) == 2)
            if (const WalkKernelSet* s = walk_kernel_set(nw, ng)) return s;
    }
    const int64_t I = c->plan.n_internal;
    // Measured on one B200 (profiles/r2_a_sweep_*.jsonl): binary trees -- 256 groups: 4x2; 128: 2x2; 64: 2x1 (7 % ahead of 1x2
    // on the 16 M-node tree, 30 % on 1 000 taxa x 5 M columns); 32: 1x1.  Trees with polytomies: one group per thread at every
    // width (the any-arity joins keep their counters in registers: half the state per thread; 4x1 is 15 % / 50 % ahead of 4x2
    // for Fitch / Sankoff on a 20 000-taxon tree with 60 % of its edges collapsed).
    static const int binary_first[4][2] = {{4, 2}, {2, 2}, {2, 1}, {1, 1}};
    static const int poly_first[3][2] = {{4, 1}, {2, 1}, {1, 1}};
    const int (*order)[2] = c->has_poly ? poly_first : binary_first;
    const int n_order = c->has_poly ? 3 : 4;
    if ((int64_t)c->tree_height * 4 > I) {  // ladder-like: tiles are the parallelism
        for (int i = 0; i < n_order; i++) {
            const WalkKernelSet* k = walk_kernel_set(order[i][0], order[i][1]);
            if (max_tile_groups > 0 && k->shape.groups() > max_tile_groups && k->shape.groups() > 32) continue;
            if ((groups + k->shape.groups() - 1) / k->shape.groups() >= 2 * 148) return k;
        }
        return walk_kernel_set(1, 1);
    }
    // Otherwise: estimated time = (cost of a tile) x (rounds of work items over the CTA slots).  A tile's cost grows
    // sub-linearly with its width and hardly shrinks when the tile is partly empty (per-op bookkeeping, program and
    // descriptor traffic do not depend on the fill): cost(tile) = c_shape * (0.7 + 0.3 * fill), c_shape relative to
    // 1x1 = 1 from the sweeps on the 16 M-node tree (Fitch 1 : 1.72 : 2.94 : 4.4, Sankoff 1 : 1.89 : 2.7 : 4.8 for
    // 1x1 : 2x1 : 2x2 : 4x2; 4x1 3.2).  With more items than slots (~5 CTAs on each of 148 SMs) the run costs the sum
    // of its tiles: 112 groups of the 16 M-node tree are one 2x2 tile (2.74) rather than four 1x1 tiles (3.85) -- a
    // rule on padding alone picked the latter and made the rank with the narrowest column shard the slowest.  With
    // fewer (a small tree over few columns) the narrowest tile that keeps the item count under the slots wins.
    auto tile_cost = [](int nw, int ng) { return nw * ng == 8 ? 4.5 : nw * ng == 4 ? (ng == 2 ? 2.85 : 3.2) : nw * ng == 2 ? 1.78 : 1.0; };
    const double chunks = std::max<double>(1.0, (double)c->plan.n_nodes / 36.0), slots = 740.0;
    const WalkKernelSet* best = walk_kernel_set(1, 1);
    double best_cost = 1e300;
    for (int i = 0; i < n_order; i++) {
        const WalkKernelSet* k = walk_kernel_set(order[i][0], order[i][1]);
        const int64_t w = k->shape.groups(), tiles = (groups + w - 1) / w;
        if (max_tile_groups > 0 && w > max_tile_groups && w > 32) continue;
        const double fill = (double)groups / (double)(w * tiles);
        const double cost = tile_cost(order[i][0], order[i][1]) * (0.7 + 0.3 * fill) * std::max(1.0, chunks * (double)tiles / slots);
        if (cost < best_cost) {
            best_cost = cost;
            best = k;
        }
    }
    return best;
}

extern "C" {

const char* pmfs_last_error(void) { return g_error.c_str(); }
const char* pmfs_version(void) { return "pmfs 0.1 (sm_100a)"; }

int pmfs_create(pmfs_ctx** out, const pmfs_tree* tree, int device) {
    if (!out || !tree) {
        g_error = "pmfs_create: null argument";
        return PMFS_ERR_ARG;
    }
    *out = nullptr;
    // validate before touching CUDA so that malformed trees are reported even without a device
    {
        Plan tmp;
        std::string err = build_plan(*tree, 1 << 30, tmp);
        if (!err.empty()) {
            g_error = "pmfs_create: " + err;
            return PMFS_ERR_ARG;
        }
    }
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0) {
        cudaGetLastError();
        g_error = "pmfs_create: no CUDA device available (this library has no CPU path)";
        return PMFS_ERR_CUDA;
    }
    if (device < 0 || device >= n_dev) {
        g_error = "pmfs_create: bad device index";
        return PMFS_ERR_ARG;
    }
    PMFS_CUDA(cudaSetDevice(device));
    pmfs_ctx* c = new pmfs_ctx();
    c->device = device;
    const int32_t V = tree->n_nodes;
    c->parent.assign(tree->parent, tree->parent + V);
    c->child_off.assign(tree->child_off, tree->child_off + V + 1);
    c->child_idx.assign(tree->child_idx, tree->child_idx + tree->child_off[V]);
    c->leaf_row.assign(tree->leaf_row, tree->leaf_row + V);
    cudaError_t se = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (se != cudaSuccess) {
        g_error = std::string("cudaStreamCreate: ") + cudaGetErrorString(se);
        delete c;
        return PMFS_ERR_CUDA;
    }
    c->own_stream = c->stream;
    {
        cudaError_t ce = cudaSuccess;
        for (auto& ev : c->ev)
            if (ce == cudaSuccess) ce = cudaEventCreate(&ev);
        for (auto& ev : c->ev_slab)
            if (ce == cudaSuccess) ce = cudaEventCreate(&ev);
        if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
        if (ce != cudaSuccess) {
            g_error = std::string("pmfs_create: event / stream creation: ") + cudaGetErrorString(ce);
            pmfs_destroy(c);
            return PMFS_ERR_CUDA;
        }
    }
    {
        std::vector<int32_t> depth((size_t)V, 0);  // internal nodes above a node, itself included
        for (int32_t v = 0; v < V; v++) {
            const bool internal = tree->child_off[v + 1] > tree->child_off[v];
            depth[v] = (v == 0 ? 0 : depth[tree->parent[v]]) + (internal ? 1 : 0);
            c->tree_height = std::max(c->tree_height, depth[v]);
        }
    }
    int rc = replan(c, auto_chunk_nodes(V, 1024, 256));
    if (rc) {
        pmfs_destroy(c);
        return rc;
    }
    c->chunk_nodes_auto = true;
    *out = c;
    return PMFS_OK;
}

void pmfs_destroy(pmfs_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    c->d_fwd_stream.release(); c->d_bwd_stream.release(); c->d_fwd_soff.release(); c->d_bwd_soff.release();
    c->d_fwd_slen.release(); c->d_bwd_slen.release(); c->d_chunk_parent.release(); c->d_chunk_nchild.release();
    c->d_sync.release(); c->d_fwd_items.release(); c->d_bwd_items.release(); c->d_rkeys_alt.release(); c->d_sort_tmp2.release();
    c->d_fwd_prog.release(); c->d_bwd_prog.release(); c->d_fwd_off.release(); c->d_bwd_off.release();
    c->d_fwd_len.release(); c->d_bwd_len.release(); c->d_fwd_order.release(); c->d_bwd_order.release();
    c->d_int_to_node.release(); c->d_row_to_node.release();
    c->d_S.release(); c->d_D.release(); c->d_cons_planes.release(); c->d_forced_planes.release();
    c->d_xcode.release(); c->d_xvalid.release(); c->d_keys.release(); c->d_keys_alt.release();
    c->d_count.release();
    c->d_scan_tmp.release(); c->d_cell_count.release(); c->d_cell_start.release(); c->d_cell_prefix.release();
    if (c->h_stats) cudaFreeHost(c->h_stats);
    c->d_sort_tmp.release(); c->d_out.release(); c->d_msa.release(); c->d_cons.release(); c->d_forced.release();
    c->d_chars.release();
    c->d_synth_node.release(); c->d_synth_pslot.release(); c->d_synth_mslot.release();
    c->d_synth_state[0].release(); c->d_synth_state[1].release();
    if (c->h_out) cudaFreeHost(c->h_out);
    if (c->h_pk_out) cudaFreeHost(c->h_pk_out);
    if (c->h_pk_off) cudaFreeHost(c->h_pk_off);
    c->d_pk_run.release(); c->d_pk_idx.release(); c->d_pk_node.release(); c->d_pk_out.release(); c->d_pk_off.release();
    c->d_pk_tmp.release();
    c->d_skeys.release(); c->d_noff.release(); c->d_cuts.release();
    if (c->h_cuts) cudaFreeHost(c->h_cuts);
    if (c->h_stage) cudaFreeHost(c->h_stage);
    for (auto& ev : c->ev)
        if (ev) cudaEventDestroy(ev);
    for (auto& ev : c->ev_slab)
        if (ev) cudaEventDestroy(ev);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

int pmfs_set_tuning(pmfs_ctx* c, int32_t chunk_nodes, int64_t tile_cols) {
    if (!c || chunk_nodes < 0 || tile_cols < 0) {
        g_error = "pmfs_set_tuning: bad argument";
        return PMFS_ERR_ARG;
    }
    c->tune_chunk_nodes = chunk_nodes;
    c->tune_tile_cols = tile_cols;
    return PMFS_OK;
}

int pmfs_set_poly_split(pmfs_ctx* c, int32_t min_arity, int32_t part_children) {
    if (!c || min_arity < 0 || part_children < 0 || part_children > kPartChildren) {
        g_error = "pmfs_set_poly_split: bad argument (at most 63 children per part)";
        return PMFS_ERR_ARG;
    }
    c->tune_split_arity = min_arity;
    c->tune_part_children = part_children ? part_children : kPartChildren;
    c->chunk_nodes_used = 0;  // the next run builds its plan anew
    c->chunk_nodes_auto = false;
    return PMFS_OK;
}

int pmfs_set_tile(pmfs_ctx* c, int32_t walking_warps, int32_t groups_per_thread) {
    if (!c || (walking_warps != 0 && !walk_kernel_set(walking_warps, groups_per_thread))) {
        g_error = "pmfs_set_tile: shapes are (walking warps x column groups per thread) 4x2, 2x2, 4x1, 1x2, 2x1, 1x1; 0 = automatic";
        return PMFS_ERR_ARG;
    }
    c->tune_nw = walking_warps;
    c->tune_ng = walking_warps ? groups_per_thread : 0;
    return PMFS_OK;
}

int pmfs_set_watchdog(pmfs_ctx* c, double seconds) {
    if (!c || seconds < 0) {
        g_error = "pmfs_set_watchdog: bad argument";
        return PMFS_ERR_ARG;
    }
    c->watchdog_ns = (long long)(seconds * 1e9);
    return PMFS_OK;
}

int pmfs_set_stream(pmfs_ctx* c, void* cuda_stream) {
    if (!c) return PMFS_ERR_ARG;
    c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
    return PMFS_OK;
}

int pmfs_get_plan_info(pmfs_ctx* c, pmfs_plan_info* info) {
    if (!c || !info) return PMFS_ERR_ARG;
    const Plan& P = c->plan;
    info->n_nodes = P.n_nodes; info->n_internal = P.n_internal; info->n_leaves = P.n_leaves;
    info->n_chunks = (int32_t)P.chunks.size();
    info->n_fwd_levels = (int32_t)P.fwd_lvl_off.size() - 1;
    info->n_bwd_levels = (int32_t)P.bwd_lvl_off.size() - 1;
    info->fwd_stack_depth = P.fwd_stack_depth; info->bwd_stack_depth = P.bwd_stack_depth;
    info->max_arity = P.max_arity;
    info->program_words = (int64_t)P.fwd_prog.size() + (int64_t)P.bwd_prog.size();
    info->edges_leaf = P.edges_leaf; info->edges_internal = P.edges_internal;
    info->n_parts = P.n_parts; info->n_set_rows = P.n_set_rows;
    return PMFS_OK;
}

int pmfs_get_timing(pmfs_ctx* c, pmfs_timing* t) {
    if (!c || !t) return PMFS_ERR_ARG;
    *t = c->timing;
    return PMFS_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------- the pipeline
// ints of synchronisation scratch a pass needs for `groups` column groups: claim counters + per-(chunk,tile) counters
static size_t sync_ints(const Plan& P, int64_t groups, int tile_groups) {
    return 64 + 2 * P.chunks.size() * (size_t)((groups + tile_groups - 1) / tile_groups);
}

// Operand ring depths of a pass (log2 of leaf-row units / set slots in flight per CTA, walk_kernels.cuh).  The 256-group
// shape keeps what BASELINE config 2 was tuned with (bottom-up 4 + 1: mostly leaves, and its walk stack and registers cap
// the resident CTAs anyway; top-down 2 + 2: under 57 KB, so 4 CTAs share an SM).  Narrower tiles have smaller units: about
// the same bytes in flight per SM need deeper rings -- but not deeper than that: more shared memory per CTA costs
// resident CTAs, and on the 16 M-node tree 16 + 4 / 8 + 8 entries ran 40 % slower than 8 + 2 / 4 + 4
// (profiles/r2_b_ring_sweep.log; a ladder, one chain of dependent nodes per column tile, does not care either way: its
// walkers wait for instructions, not for operands).  PMFS_RINGS="fl,fs,bl,bs" (environment, log2 values) overrides.
static void ring_depths(const pmfs_ctx* c, const WalkKernelSet& ks, bool fwd, int algo, int stack_depth, int32_t* leaf_log2, int32_t* set_log2) {
    const int groups = ks.shape.groups();
    int fl, fs, bl, bs;
    if (groups >= 256) { fl = 2; fs = 0; bl = 1; bs = 1; }
    else if (groups >= 128) { fl = 2; fs = 1; bl = 2; bs = 1; }
    else if (groups >= 64) { fl = 3; fs = 1; bl = 2; bs = 2; }
    else { fl = 3; fs = 2; bl = 3; bs = 2; }
    // The 2x1 top-down Fitch kernel is compiled for 64 registers (walk_kernels.cuh bwd_min_ctas): 10 CTAs per SM by
    // registers, which shared memory only lets in with a two-slot set ring (8.88 -> 8.40 ms per 2 048-column batch of the
    // 16 M-node tree; with the four-slot ring 64 registers gain nothing, profiles/r2_x_regs64.log)
    if (algo == PMFS_FITCH && ks.shape.nw == 2 && ks.shape.ng == 1) bs = 1;
    if (const char* e = getenv("PMFS_RINGS")) {
        int a, b, d, f;
        if (sscanf(e, "%d,%d,%d,%d", &a, &b, &d, &f) == 4) { fl = a; fs = b; bl = d; bs = f; }
    }
    auto clamp = [](int v, int hi) { return v < 0 ? 0 : (v > hi ? hi : v); };
    int l = clamp(fwd ? fl : bl, 5), t = clamp(fwd ? fs : bs, 4);  // kMaxLeafRing = 32, kMaxSetRing = 16
    // (whatever was asked for: the rings and the walk stack have to fit in one CTA's shared memory)
    const size_t stack = (size_t)stack_depth * (fwd ? ks.fwd_smem_per_level : ks.bwd_smem_per_level);
    while (ks.smem_fixed + (ks.leaf_unit_bytes << l) + (ks.set_slot_bytes << t) + stack > 160 * 1024 && (l > 0 || t > 0)) {
        if ((ks.leaf_unit_bytes << l) >= (ks.set_slot_bytes << t) && l > 0) l--;
        else if (t > 0) t--;
        else l--;
    }
    *leaf_log2 = l;
    *set_log2 = t;
}

// One launch per pass over all chunks (the name is historical: the first version launched once per chunk level).
// wp.ready must point to sync_ints() ints of device memory.
template <int ALGO>
static int launch_levels(pmfs_ctx* c, WalkParams wp, bool fwd, int64_t* launches, bool sync_cleared = false) {
    const Plan& P = c->plan;
    wp.prog = fwd ? c->d_fwd_prog.ptr : c->d_bwd_prog.ptr;
    wp.chunk_off = fwd ? c->d_fwd_off.ptr : c->d_bwd_off.ptr;
    wp.chunk_len = fwd ? c->d_fwd_len.ptr : c->d_bwd_len.ptr;
    wp.stream = fwd ? c->d_fwd_stream.ptr : c->d_bwd_stream.ptr;
    wp.stream_off = fwd ? c->d_fwd_soff.ptr : c->d_bwd_soff.ptr;
    wp.stream_len = fwd ? c->d_fwd_slen.ptr : c->d_bwd_slen.ptr;
    wp.order = fwd ? c->d_fwd_order.ptr : c->d_bwd_order.ptr;
    wp.items = fwd ? c->d_fwd_items.ptr : c->d_bwd_items.ptr;
    wp.chunk_parent = c->d_chunk_parent.ptr;
    wp.chunk_nchild = c->d_chunk_nchild.ptr;
    const WalkKernelSet& ks = *c->ks;
    const int tile_groups = ks.shape.groups(), threads = ks.shape.threads();
    wp.ctas_per_chunk = (int32_t)((wp.G + tile_groups - 1) / tile_groups);
    wp.stack_depth = fwd ? P.fwd_stack_depth : P.bwd_stack_depth;
    wp.spin_limit_ns = c->watchdog_ns;
    {
        static const int lead = getenv("PMFS_CLAIM_LEAD") ? atoi(getenv("PMFS_CLAIM_LEAD")) : 4;
        wp.claim_lead = lead;
    }
    const int64_t n_items = (int64_t)P.chunks.size() * wp.ctas_per_chunk;
    if (n_items > 0x7fffffffll) {
        g_error = "too many (chunk, column tile) work items: use a smaller column tile";
        return PMFS_ERR_ARG;
    }
    wp.n_items = (int32_t)n_items;
    int* sync = wp.ready;
    const size_t per_pass = (size_t)n_items;
    wp.next_item = sync + (fwd ? 0 : 32);
    wp.ready = sync + 64 + (fwd ? 0 : per_pass);
    ring_depths(c, ks, fwd, ALGO, fwd ? P.fwd_stack_depth : P.bwd_stack_depth, &wp.leaf_ring_log2, &wp.set_ring_log2);
    const size_t smem = ks.smem_fixed + (ks.leaf_unit_bytes << wp.leaf_ring_log2) + (ks.set_slot_bytes << wp.set_ring_log2) +
                        (fwd ? (size_t)P.fwd_stack_depth * ks.fwd_smem_per_level : (size_t)P.bwd_stack_depth * ks.bwd_smem_per_level);
    if (smem > 200 * 1024) {
        g_error = "walk stack does not fit in shared memory (tree too unbalanced for this chunk size)";
        return PMFS_ERR_ARG;
    }
    const void* k = fwd ? ks.fwd[ALGO][c->has_poly ? 1 : 0] : ks.bwd[ALGO];
    int per_sm = 0;
    for (auto& e : c->occ_cache)
        if (e.k == k && e.smem == smem) per_sm = e.per_sm;
    if (per_sm == 0) {
        size_t allowed = 40 * 1024;  // the attribute is a maximum: only ever raise it
        for (auto& e : c->occ_cache)
            if (e.k == k) allowed = std::max(allowed, e.smem);
        if (smem > allowed) PMFS_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        PMFS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, threads, smem));
        if (per_sm < 1) per_sm = 1;
        c->occ_cache.push_back({k, smem, per_sm});
    }
    if (c->n_sms == 0) PMFS_CUDA(cudaDeviceGetAttribute(&c->n_sms, cudaDevAttrMultiProcessorCount, c->device));
    const unsigned grid = (unsigned)std::min<int64_t>(n_items, (int64_t)c->n_sms * per_sm);
    // claim counters and dependency counters of BOTH passes are cleared once, ahead of the bottom-up pass
    if (fwd && !sync_cleared) PMFS_CUDA(cudaMemsetAsync(sync, 0, (64 + 2 * per_pass) * sizeof(int), c->stream));
    void* args[] = {&wp};
    if (ks.prof_reset) {  // -DPMFS_WAITPROF builds only
        PMFS_CUDA(cudaStreamSynchronize(c->stream));
        ks.prof_reset();
    }
    PMFS_CUDA(cudaLaunchKernel(k, dim3(grid), dim3(threads), args, smem, c->stream));
    if (ks.prof_dump) {
        PMFS_CUDA(cudaStreamSynchronize(c->stream));
        ks.prof_dump(fwd, grid, per_sm, wp.n_items);
    }
    (*launches)++;
    return PMFS_OK;
}

static int launch_order_cells(pmfs_ctx* c, size_t region_cap, int64_t cells, int64_t W, int pos_bits, const unsigned long long* d_stats,
                              size_t out_cap) {
    const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>((int64_t)c->n_sms * 8, (cells + 255) / 256));
    if (c->keys_out)
        order_cells_kernel<true><<<grid, 256, 0, c->stream>>>(c->d_keys.ptr, region_cap, c->d_cell_count.ptr, c->d_cell_start.ptr,
                                                              c->d_cell_prefix.ptr, cells, (int32_t)W, pos_bits, d_stats, out_cap, c->d_skeys.ptr);
    else
        order_cells_kernel<false><<<grid, 256, 0, c->stream>>>(c->d_keys.ptr, region_cap, c->d_cell_count.ptr, c->d_cell_start.ptr,
                                                               c->d_cell_prefix.ptr, cells, (int32_t)W, pos_bits, d_stats, out_cap, c->d_out.ptr);
    PMFS_CUDA(cudaGetLastError());
    return PMFS_OK;
}

// Column groups whose set arrays (S, and D1 for Sankoff) fit in the device memory that is free now (what the context
// already holds for them counts as free), with `extra_per_group` bytes per group for whatever else the caller allocates
// per column group, and 30 % left aside for records and cells (on a 16 M-node tree the record buffers of a column
// group -- emitted, ordered, received, converted -- come to a fifth of its set arrays).
namespace pmfs {
int mem_limited_groups(pmfs_ctx* c, int algo, int64_t G_all, size_t extra_per_group, int64_t* groups) {
    const int64_t I = c->plan.n_set_rows;
    const int64_t per_group = I * 64 * (algo == PMFS_SANKOFF ? 2 : 1) + 64 + (int64_t)extra_per_group;
    const size_t need = (size_t)I * 4 * (size_t)G_all;
    *groups = G_all;
    if (extra_per_group == 0 && c->d_S.cap >= need && (algo != PMFS_SANKOFF || c->d_D.cap >= need)) return PMFS_OK;
    // (cudaMemGetInfo costs a fraction of a millisecond: only asked when buffers may have to grow)
    size_t free_b = 0, total_b = 0;
    PMFS_CUDA(cudaMemGetInfo(&free_b, &total_b));
    free_b += c->d_S.cap * sizeof(uint4) + c->d_D.cap * sizeof(uint4);  // reusable
    *groups = std::min<int64_t>(G_all, std::max<int64_t>(1, (int64_t)(free_b * 0.7) / per_group));
    return PMFS_OK;
}

}  // namespace pmfs

namespace pmfs {
int run_device_impl(pmfs_ctx* c, int algo, const uint8_t* d_msa, int64_t row_stride, int32_t n_rows, int64_t n_cols,
                    const uint8_t* d_cons, const uint8_t* d_forced, pmfs_mut** d_out, int64_t* n_out) {
    if (!c || (algo != PMFS_FITCH && algo != PMFS_SANKOFF) || !d_msa || !d_cons || !n_out || n_cols <= 0) {
        g_error = "pmfs_run: bad argument";
        return PMFS_ERR_ARG;
    }
    if (row_stride % 16 != 0 || ((uintptr_t)d_msa & 15) != 0 || row_stride * 2 < n_cols) {
        g_error = "pmfs_run_device: MSA rows must be 16-byte aligned with a stride that is a multiple of 16 and covers n_cols";
        return PMFS_ERR_ARG;
    }
    if ((int64_t)c->plan.row_to_node.size() > (int64_t)n_rows) {
        g_error = "pmfs_run: a leaf refers to an MSA row >= n_rows";
        return PMFS_ERR_ARG;
    }
    if (n_cols > 0xffffffffll) {
        g_error = "pmfs_run: more than 2^32 columns";
        return PMFS_ERR_ARG;
    }
    PMFS_CUDA(cudaSetDevice(c->device));
    c->too_many_records = false;
    const auto host_t0 = std::chrono::steady_clock::now();
    if (c->plan.n_parts > 0 && c->chunk_nodes_used > 0 && c->plan.part_rows != (algo == PMFS_SANKOFF ? kPartRowsSankoff : kPartRowsFitch)) {
        int rc = replan(c, c->chunk_nodes_used, algo);  // split polytomies: the other algorithm's partial results
        if (rc) return rc;
    }
    const Plan* P = &c->plan;
    const int64_t G_all = (n_cols + 31) / 32;
    const int64_t I = P->n_internal;

    // column tile so that the set arrays fit in HBM
    int64_t tile_groups = G_all;
    auto set_elems = [&](int64_t groups) {  // uint4 elements of one tile-major set array for `groups` column groups
        return (size_t)c->plan.n_set_rows * 4 * (size_t)groups;  // (n_set_rows does not depend on the chunk size)
    };
    {
        int rc = mem_limited_groups(c, algo, G_all, 0, &tile_groups);
        if (rc) return rc;
    }
    if (c->tune_tile_cols > 0) tile_groups = std::min<int64_t>(tile_groups, (c->tune_tile_cols + 31) / 32);
    tile_groups = std::min(tile_groups, G_all);
    c->ks = choose_shape(c, tile_groups, tile_groups < G_all ? tile_groups : 0);
    const int TG = c->ks->shape.groups();  // column groups per work item
    if (tile_groups < G_all) tile_groups = std::max<int64_t>(TG, tile_groups / TG * TG);
    tile_groups = std::min(tile_groups, G_all);

    const int32_t want_chunk = c->tune_chunk_nodes > 0 ? c->tune_chunk_nodes : auto_chunk_nodes(P->n_nodes, tile_groups, TG);
    // (an automatic target within 4x of the plan in use keeps that plan: column batches of different widths -- the ragged
    // last batch of a run -- would otherwise re-plan back and forth, seconds each on a 16 M-node tree, for chunk sizes
    // that the walk-stack cap decides anyway)
    const bool close_enough = c->tune_chunk_nodes == 0 && c->chunk_nodes_used > 0 && want_chunk <= 4 * (int64_t)c->chunk_nodes_used &&
                              c->chunk_nodes_used <= 4 * (int64_t)want_chunk && c->chunk_nodes_auto;
    if (want_chunk != c->chunk_nodes_used && !close_enough) {
        int rc = replan(c, want_chunk, algo);
        if (rc) return rc;
        c->chunk_nodes_auto = c->tune_chunk_nodes == 0;
        P = &c->plan;
    }
    const int64_t NC = (int64_t)P->chunks.size();
    int pos_bits = 1, node_bits = 1;
    const int64_t pos_span = c->shard_total_cols > 0 ? c->shard_total_cols : n_cols;
    if (c->shard_total_cols > 0 && c->shard_pos_offset + n_cols > c->shard_total_cols) {
        g_error = "pmfs_run: column shard exceeds the total column count given to pmfs_set_shard";
        return PMFS_ERR_ARG;
    }
    while ((1ll << pos_bits) < pos_span) pos_bits++;
    while ((1ll << node_bits) < P->n_nodes) node_bits++;
    if (pos_bits + node_bits + kPayloadBits > 64) {
        g_error = "pmfs_run: log2(nodes) + log2(columns) must not exceed 54 (a record is one 64-bit word)";
        return PMFS_ERR_ARG;
    }

    int rc;
    if ((rc = c->d_S.ensure(set_elems(tile_groups)))) return rc;
    if (algo == PMFS_SANKOFF && (rc = c->d_D.ensure(set_elems(tile_groups)))) return rc;
    if ((rc = c->d_cons_planes.ensure(tile_groups))) return rc;
    if (d_forced && (rc = c->d_forced_planes.ensure(tile_groups))) return rc;
    if ((rc = c->d_xcode.ensure((size_t)NC * tile_groups))) return rc;
    if ((rc = c->d_xvalid.ensure((size_t)NC * tile_groups))) return rc;
    if ((rc = c->d_sync.ensure(sync_ints(*P, tile_groups, TG)))) return rc;
    // output regions: per owning rank in shard mode, else one per 32 column groups (see DevEnv::emit)
    const int64_t W = (G_all + 31) / 32;
    const int n_parts = std::max(1, c->parts.n_parts);
    const bool shard = c->parts.n_parts > 0;
    const bool by_columns = shard ? W * n_parts <= 65535 : W <= (1 << 20);
    // single-GPU runs: low node bits that select the output bucket (see DevEnv::emit) -- as few as make the remaining
    // node bits fit the smallest number of 8-bit radix passes, with at most 2^20 output areas
    int bucket_bits = 0;
    if (!shard && by_columns) {
        int max_bits = 10;
        while (max_bits > 0 && (W << max_bits) > (1 << 20)) max_bits--;
        max_bits = std::min(max_bits, node_bits);
        const int passes = (node_bits - max_bits + 7) / 8;
        bucket_bits = std::max(0, node_bits - 8 * passes);
    }
    // ... unless the (node, region) cell matrix fits: then the records are ordered without any sort (order_cells_kernel)
    static const bool force_radix = getenv("PMFS_ORDER") && std::string(getenv("PMFS_ORDER")) == "radix";  // test hook
    const int64_t cells = (int64_t)P->n_nodes * W;
    const bool use_cells = !shard && by_columns && !force_radix && cells <= (1ll << 30);
    if (use_cells) {
        bucket_bits = 0;
        if ((rc = c->d_cell_count.ensure((size_t)cells))) return rc;
        if ((rc = c->d_cell_start.ensure((size_t)cells))) return rc;
        if ((rc = c->d_cell_prefix.ensure((size_t)cells))) return rc;
    }
    const int64_t n_owners = !by_columns ? n_parts : shard ? W * n_parts : (W << bucket_bits);  // number of output areas
    const size_t n_ctr = (size_t)std::max<int64_t>(16, n_owners);
    if ((rc = c->d_count.ensure(2 * n_ctr + 16))) return rc;
    // [n_ctr area counters][16: sum, max of the counters, error flag][n_ctr area offsets]
    unsigned long long* const d_stats = c->d_count.ptr + n_ctr;
    unsigned long long* const d_off = d_stats + 16;
    if (!c->h_stats) PMFS_CUDA(cudaMallocHost(&c->h_stats, 16 * sizeof(unsigned long long)));
    size_t cap = std::max<size_t>(c->d_keys.cap, 1 << 20);  // total records; one region is cap / n_owners
    if (use_cells && c->d_keys.cap == 0) {
        // First call: every (chunk, column region) reserves output blocks of 256 records (DevEnv::emit_cells); two
        // blocks each covers the usual mutation densities, so that the passes need not be run twice to find the size.
        // Bounded by what is free on the device; denser data still takes the grow-and-redo path below.
        size_t free_b = 0, total_b = 0;
        cudaMemGetInfo(&free_b, &total_b);
        const size_t guess = (size_t)NC * (size_t)W * 512;
        if (guess * sizeof(unsigned long long) < free_b / 4) cap = std::max(cap, guess);
    }

    pmfs_timing& tm = c->timing;
    const float h2d_keep = tm.h2d_ms;
    tm = pmfs_timing();
    tm.h2d_ms = h2d_keep;
    float fwd_ms = 0, bwd_ms = 0;
    unsigned long long count = 0;
    std::vector<unsigned long long>& counts = c->h_counts;
    counts.assign(shard ? n_ctr : 16, 0);
    size_t region = 0, ordered_cap = 0;
    for (int attempt = 0; attempt < 2; attempt++) {
        if ((rc = c->d_keys.ensure(cap))) return rc;
        region = c->d_keys.cap / (size_t)n_owners;
        // (the area counters and the walk kernels' claim / dependency counters are cleared by the prep kernel below; the
        // cell matrix is left clear by order_cells_kernel and only needs a memset after a run that did not get there)
        if (use_cells && c->cells_clean_n != cells)
            PMFS_CUDA(cudaMemsetAsync(c->d_cell_count.ptr, 0, (size_t)cells * sizeof(uint32_t), c->stream));
        c->cells_clean_n = 0;
        fwd_ms = bwd_ms = 0;
        tm.n_launches = 0;
        // column segments: the memory tiles, cut once more where pmfs_run's second upload slab begins
        const int64_t split = c->slab_split > 0 && c->slab_split < G_all ? c->slab_split : 0;
        const bool multi = tile_groups < G_all || split > 0;
        for (int64_t g0 = 0, g1; g0 < G_all; g0 = g1) {
            g1 = std::min(g0 + tile_groups, G_all);
            if (g0 < split && split < g1) g1 = split;
            if (split > 0 && (g0 == 0 || g0 == split))  // the slab's bytes have arrived (copy stream of pmfs_run)
                PMFS_CUDA(cudaStreamWaitEvent(c->stream, c->ev_slab[g0 == 0 ? 0 : 1], 0));
            WalkParams wp = {};
            wp.G = g1 - g0;
            wp.col_base = g0 * 32;
            wp.tile_cols = std::min<int64_t>(wp.G * 32, n_cols - wp.col_base);
            wp.msa = (const uint4*)(d_msa + g0 * 16);
            wp.msa_stride16 = row_stride / 16;
            wp.S = c->d_S.ptr;
            wp.D = c->d_D.ptr;
            wp.cons_planes = c->d_cons_planes.ptr;
            wp.forced_planes = d_forced ? c->d_forced_planes.ptr : nullptr;
            wp.xcode = c->d_xcode.ptr;
            wp.xvalid = c->d_xvalid.ptr;
            wp.int_to_node = c->d_int_to_node.ptr;
            wp.row_to_node = c->d_row_to_node.ptr;
            wp.out_keys = c->d_keys.ptr;
            wp.out_count = c->d_count.ptr;
            wp.out_cap = region;
            wp.n_owners = n_parts;
            wp.n_regions = by_columns ? (int32_t)W : 0;
            wp.bucket_mask = (1u << bucket_bits) - 1u;
            wp.cell_count = use_cells ? c->d_cell_count.ptr : nullptr;
            wp.cell_start = use_cells ? c->d_cell_start.ptr : nullptr;
            wp.region_base = g0 / 32;
            for (int r = 0; r <= c->parts.n_parts; r++) wp.owner_bound[r] = c->parts.bound[r];
            wp.err_flag = (unsigned int*)(d_stats + 2);
            wp.pos_bits = pos_bits;
            wp.pos_offset = c->shard_total_cols > 0 ? c->shard_pos_offset : 0;
            wp.ready = c->d_sync.ptr;
            const unsigned pg = (unsigned)((wp.G + 127) / 128);
            PMFS_CUDA(cudaEventRecord(c->ev[0], c->stream));
            {
                const int64_t seg_items = (int64_t)P->chunks.size() * ((wp.G + TG - 1) / TG);
                prep_planes_kernel<<<pg, 128, 0, c->stream>>>(d_cons + wp.col_base, wp.tile_cols, wp.G, c->d_cons_planes.ptr,
                                                             (uint32_t*)c->d_sync.ptr, 64 + 2 * seg_items,
                                                             g0 == 0 ? (uint32_t*)c->d_count.ptr : nullptr,
                                                             g0 == 0 ? (int64_t)(n_ctr + 16) * 2 : 0);
            }
            tm.n_launches++;
            if (d_forced) {
                prep_planes_kernel<<<pg, 128, 0, c->stream>>>(d_forced + wp.col_base, wp.tile_cols, wp.G, c->d_forced_planes.ptr);
                tm.n_launches++;
            }
            rc = algo == PMFS_FITCH ? launch_levels<kAlgoFitch>(c, wp, true, &tm.n_launches, true)
                                    : launch_levels<kAlgoSankoff>(c, wp, true, &tm.n_launches, true);
            if (rc) return rc;
            PMFS_CUDA(cudaEventRecord(c->ev[1], c->stream));
            rc = algo == PMFS_FITCH ? launch_levels<kAlgoFitch>(c, wp, false, &tm.n_launches)
                                    : launch_levels<kAlgoSankoff>(c, wp, false, &tm.n_launches);
            if (rc) return rc;
            PMFS_CUDA(cudaEventRecord(c->ev[2], c->stream));
            if (multi) {  // several segments reuse the events: read them before the next one records again
                PMFS_CUDA(cudaEventSynchronize(c->ev[2]));
                float a = 0, b = 0;
                cudaEventElapsedTime(&a, c->ev[0], c->ev[1]);
                cudaEventElapsedTime(&b, c->ev[1], c->ev[2]);
                fwd_ms += a;
                bwd_ms += b;
            }
        }
        PMFS_CUDA(cudaEventRecord(c->ev[3], c->stream));
        unsigned long long* const stats = c->h_stats;
        if (shard) {
            PMFS_CUDA(cudaMemcpyAsync(counts.data(), c->d_count.ptr, counts.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
            PMFS_CUDA(cudaMemcpyAsync(stats, d_stats, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
        } else {
            area_stats_kernel<<<(unsigned)std::min<int64_t>(296, (n_owners + 255) / 256), 256, 0, c->stream>>>(c->d_count.ptr, n_owners, d_stats);
            tm.n_launches++;
            if (!use_cells) PMFS_CUDA(cudaMemcpyAsync(stats, d_stats, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
            if (use_cells) {
                // exclusive prefix sum over the cell counts = every cell's place in the (node,pos) order; the copy into
                // that order is queued right away when the output buffer of the previous run is still there (the kernel
                // checks on the device that nothing overflowed), so the GPU does not wait for the host to learn the count
                size_t scan_bytes = 0;
                PMFS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, c->d_cell_count.ptr, c->d_cell_prefix.ptr, (int)cells, c->stream));
                if ((rc = c->d_scan_tmp.ensure(scan_bytes + 16))) return rc;
                PMFS_CUDA(cub::DeviceScan::ExclusiveSum(c->d_scan_tmp.ptr, scan_bytes, c->d_cell_count.ptr, c->d_cell_prefix.ptr, (int)cells, c->stream));
                // (the region counters include the unused ends of the warps' output blocks: the record count is the cells')
                cells_total_kernel<<<1, 1, 0, c->stream>>>(c->d_cell_count.ptr, c->d_cell_prefix.ptr, cells, d_stats);
                tm.n_launches++;
                PMFS_CUDA(cudaMemcpyAsync(stats, d_stats, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
                if (c->keys_out) {  // per-node offsets and the owners' cut points only need the prefix sums: queue them now
                    if ((rc = c->d_noff.ensure((size_t)P->n_nodes + 1))) return rc;
                    if ((rc = c->d_cuts.ensure(32))) return rc;
                    if (!c->h_cuts) PMFS_CUDA(cudaMallocHost(&c->h_cuts, 32 * sizeof(unsigned long long)));
                    node_offsets_kernel<<<(unsigned)((P->n_nodes + 256) / 256), 256, 0, c->stream>>>(c->d_cell_prefix.ptr, (int32_t)W, P->n_nodes, d_stats,
                                                                                                   c->d_noff.ptr, c->xparts, c->d_cuts.ptr);
                    tm.n_launches++;
                    PMFS_CUDA(cudaMemcpyAsync(c->h_cuts, c->d_cuts.ptr, 17 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
                }
                ordered_cap = c->keys_out ? c->d_skeys.cap : c->d_out.cap;
                if (ordered_cap > 0) {
                    rc = launch_order_cells(c, region, cells, W, pos_bits, d_stats, ordered_cap);
                    if (rc) return rc;
                    tm.n_launches++;
                    PMFS_CUDA(cudaEventRecord(c->ev[7], c->stream));  // end of the step if nothing has to be redone
                }
            } else if (by_columns) {
                // Areas are laid out (low node bits, column region)-major and every node's records inside an area are
                // contiguous and column-ordered: pack the areas; a stable sort on the remaining HIGH node bits alone then
                // gives the (node,pos) order.  Queued before the host knows the record count (the device checks for
                // overflow itself), so the GPU packs while the host waits for the count the sort call needs.
                size_t scan_bytes = 0;
                PMFS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, c->d_count.ptr, d_off, (int)n_owners, c->stream));
                if ((rc = c->d_scan_tmp.ensure(scan_bytes + 16))) return rc;
                if ((rc = c->d_keys_alt.ensure(c->d_keys.cap))) return rc;
                PMFS_CUDA(cub::DeviceScan::ExclusiveSum(c->d_scan_tmp.ptr, scan_bytes, c->d_count.ptr, d_off, (int)n_owners, c->stream));
                const unsigned cg = (unsigned)std::max<int64_t>(1, std::min<int64_t>(148 * 8, (n_owners + 7) / 8));
                compact_areas_kernel<<<cg, 256, 0, c->stream>>>(c->d_keys.ptr, region, c->d_count.ptr, d_off, n_owners, d_stats + 1,
                                                                c->d_keys_alt.ptr);
                tm.n_launches++;
            }
        }
        PMFS_CUDA(cudaStreamSynchronize(c->stream));
        const unsigned int err = (unsigned int)stats[2];
        if (use_cells && stats[3] >> 32) {
            // the cells' prefix sums are 32-bit: with 2^32 or more record slots reserved they could wrap (the ordering
            // kernel has refused to run on the device for the same reason)
            c->too_many_records = true;
            g_error = "pmfs_run: too many records for one call (2^32 record slots reserved): run fewer columns per call -- "
                      "pmfs_run_batched does that";
            return PMFS_ERR_ARG;
        }
        if (!multi) {
            cudaEventElapsedTime(&fwd_ms, c->ev[0], c->ev[1]);
            cudaEventElapsedTime(&bwd_ms, c->ev[1], c->ev[2]);
        }
        if (err) {
            g_error = "pmfs_run: Sankoff root has no finite cost in some column and no forced root was given "
                      "(the reference asserts here, src/fitchSankoff.cpp:505)";
            return PMFS_ERR_ROOT_INVALID;
        }
        count = shard ? 0 : stats[0];
        unsigned long long worst = shard ? 0 : stats[1];
        for (int64_t r = 0; shard && r < n_owners; r++) {
            count += counts[r];
            worst = std::max(worst, counts[r]);
        }
        if (worst <= region) break;
        if (attempt == 1) {
            g_error = "pmfs_run: record buffer overflow after regrowth";
            return PMFS_ERR_NOMEM;
        }
        cap = (size_t)n_owners * ((size_t)worst + (size_t)worst / 4 + 64);  // too small: grow and redo the passes
    }
    tm.forward_ms = fwd_ms;
    tm.backward_ms = bwd_ms;
    {
        const int64_t Gb = G_all * 16;  // bytes per node row
        const int64_t sets = (algo == PMFS_SANKOFF) ? 2 : 1;
        tm.forward_bytes = P->n_rows_used * Gb + I * Gb * 4 * sets;
        tm.backward_bytes = P->n_rows_used * Gb + I * Gb * 4;
    }

    const int64_t n = (int64_t)count;
    if (n > 0x7fffffffll) {
        c->too_many_records = true;
        g_error = "pmfs_run: more than 2^31 records in one call: run fewer columns per call -- pmfs_run_batched does that";
        return PMFS_ERR_ARG;
    }
    if (c->parts.n_parts > 0) {
        // shard mode: records were emitted straight into one region per owning rank and stay unordered; the owner
        // orders what it receives (pmfs_sort_records)
        c->parts_by_columns = by_columns;
        if (by_columns) {
            // pack the (owner, column region) regions: every owner's records become one contiguous block whose regions
            // are in column order, which is what lets the owner order its inbox with a stable sort on the node bits
            if ((rc = c->d_keys_alt.ensure(c->d_keys.cap))) return rc;
            std::vector<unsigned long long>& off = c->h_offsets;
            off.assign(counts.size(), 0);
            for (int r = 1; r < n_owners; r++) off[r] = off[r - 1] + counts[r - 1];
            for (int o = 0; o < 16; o++) {
                c->part_counts[o] = 0;
                c->part_offsets[o] = o < n_parts ? (int64_t)off[(size_t)o * W] : n;
                for (int64_t w = 0; o < n_parts && w < W; w++) c->part_counts[o] += (int64_t)counts[(size_t)o * W + w];
            }
            if (n > 0) {
                PMFS_CUDA(cudaMemcpyAsync(d_off, off.data(), (size_t)n_owners * sizeof(unsigned long long), cudaMemcpyHostToDevice, c->stream));
                const unsigned bx = (unsigned)std::max<size_t>(1, std::min<size_t>(64, ((size_t)n / n_owners + 255) / 256));
                compact_regions_kernel<<<dim3(bx, (unsigned)n_owners), 256, 0, c->stream>>>(c->d_keys.ptr, region, c->d_count.ptr, d_off,
                                                                                           c->d_keys_alt.ptr);
                PMFS_CUDA(cudaStreamSynchronize(c->stream));
                tm.n_launches += 1;
            }
            c->parts_keys = c->d_keys_alt.ptr;
        } else {
            for (int o = 0; o < 16; o++) {
                c->part_counts[o] = o < n_parts ? (int64_t)counts[o] : 0;
                c->part_offsets[o] = (int64_t)region * o;
            }
            c->parts_keys = c->d_keys.ptr;
        }
        tm.sort_ms = 0;
        tm.total_ms = tm.h2d_ms + tm.forward_ms + tm.backward_ms;
        c->last_n_out = 0;
        tm.wall_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - host_t0).count();
        *n_out = n;
        if (d_out) *d_out = nullptr;
        return PMFS_OK;
    }
    if (n > 0 && use_cells) {
        if ((size_t)n > ordered_cap) {  // first run, or more records than ever before: now the count is known
            if (c->keys_out) {
                if ((rc = c->d_skeys.ensure((size_t)n + (size_t)n / 8))) return rc;
            } else {
                if ((rc = c->d_out.ensure((size_t)n + (size_t)n / 8))) return rc;
            }
            rc = launch_order_cells(c, region, cells, W, pos_bits, d_stats, c->keys_out ? c->d_skeys.cap : c->d_out.cap);
            if (rc) return rc;
            tm.n_launches++;
        }
    } else if (n > 0) {  // order by (node,pos): keys-only radix sort over exactly the key bits that can be set
        if ((rc = c->d_keys_alt.ensure(c->d_keys.cap))) return rc;
        if (!c->keys_out && (rc = c->d_out.ensure((size_t)n))) return rc;
        cub::DoubleBuffer<unsigned long long> kb(c->d_keys.ptr, c->d_keys_alt.ptr);
        const int lo = by_columns ? kPayloadBits + pos_bits + bucket_bits : kPayloadBits;
        const int hi = kPayloadBits + pos_bits + node_bits;
        size_t tmp_bytes = 0;
        if (lo < hi) PMFS_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, kb, (int)n, lo, hi, c->stream));
        if ((rc = c->d_sort_tmp.ensure(tmp_bytes + 16))) return rc;
        // (the areas were packed into d_keys_alt above; no node bits are left to sort on when the tree has at most
        // 2^bucket_bits nodes)
        if (by_columns) kb = cub::DoubleBuffer<unsigned long long>(c->d_keys_alt.ptr, c->d_keys.ptr);
        if (lo < hi) PMFS_CUDA(cub::DeviceRadixSort::SortKeys(c->d_sort_tmp.ptr, tmp_bytes, kb, (int)n, lo, hi, c->stream));
        if (c->keys_out) {
            c->skeys_view = kb.Current();  // the ordered 64-bit records, in one of the two sort buffers
        } else {
            assemble_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(kb.Current(), n, pos_bits, c->d_out.ptr);
        }
        PMFS_CUDA(cudaGetLastError());
        tm.n_launches += 1;
    }
    if (c->keys_out && !use_cells) {  // per-node offsets by binary search in the ordered records
        if ((rc = c->d_noff.ensure((size_t)P->n_nodes + 1))) return rc;
        if ((rc = c->d_cuts.ensure(32))) return rc;
        if (!c->h_cuts) PMFS_CUDA(cudaMallocHost(&c->h_cuts, 32 * sizeof(unsigned long long)));
        node_lower_bound_kernel<<<(unsigned)((P->n_nodes + 256) / 256), 256, 0, c->stream>>>(c->skeys_view, n, kPayloadBits + pos_bits, P->n_nodes,
                                                                                           c->d_noff.ptr, c->xparts, c->d_cuts.ptr);
        PMFS_CUDA(cudaMemcpyAsync(c->h_cuts, c->d_cuts.ptr, 17 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    } else if (c->keys_out) {
        c->skeys_view = c->d_skeys.ptr;
    }
    if (use_cells) c->cells_clean_n = cells;  // order_cells_kernel ran to the end (or nothing was emitted): the matrix is clear
    if (use_cells && ordered_cap > 0 && (size_t)n <= ordered_cap) {
        // the ordering was queued before the one synchronisation of the step and has finished with it
        cudaEventElapsedTime(&tm.sort_ms, c->ev[3], c->ev[7]);
    } else {
        PMFS_CUDA(cudaEventRecord(c->ev[4], c->stream));
        PMFS_CUDA(cudaEventSynchronize(c->ev[4]));
        cudaEventElapsedTime(&tm.sort_ms, c->ev[3], c->ev[4]);
    }
    tm.total_ms = tm.h2d_ms + tm.forward_ms + tm.backward_ms + tm.sort_ms;
    c->last_n_out = n;
    tm.wall_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - host_t0).count();
    *n_out = n;
    if (d_out) *d_out = c->d_out.ptr;
    return PMFS_OK;
}
}  // namespace pmfs

extern "C" {

int pmfs_run_device(pmfs_ctx* c, int algo, const uint8_t* d_msa, int64_t row_stride, int32_t n_rows, int64_t n_cols,
                    const uint8_t* d_cons, const uint8_t* d_forced, pmfs_mut** d_out, int64_t* n_out) {
    if (c) c->timing.h2d_ms = 0;
    return run_device_impl(c, algo, d_msa, row_stride, n_rows, n_cols, d_cons, d_forced, d_out, n_out);
}

int pmfs_set_batch_cols(pmfs_ctx* c, int64_t batch_cols) {
    if (!c || batch_cols < 0) {
        g_error = "pmfs_set_batch_cols: bad argument";
        return PMFS_ERR_ARG;
    }
    c->tune_batch_cols = batch_cols;
    return PMFS_OK;
}

}  // extern "C"

// Equal batches of at most gmax column groups whose size is a multiple of a wide tile shape (128 groups when there is room
// for two of them, else 64, else 32): a batch of, say, 157 groups would run on the narrowest shape or idle a fifth of a
// wider one.
namespace pmfs {
int64_t balanced_batch_groups(int64_t G_all, int64_t gmax) {
    if (gmax >= G_all) return G_all;
    const int64_t q = gmax >= 256 ? 128 : (gmax >= 64 ? 64 : (gmax >= 32 ? 32 : 1));
    const int64_t n_b = (G_all + gmax - 1) / gmax;
    int64_t gb = ((G_all + n_b - 1) / n_b + q - 1) / q * q;
    if (gb > gmax) gb = gmax / q * q;
    return std::max<int64_t>(1, gb);
}
}  // namespace pmfs

extern "C" {

// Columns per batch: what the caller asked for, else as many as the set arrays of one batch fit in free device memory,
// in equal batches whose size is a multiple of 32 column groups (whole output regions and tile shapes).
static int batch_groups_for(pmfs_ctx* c, int algo, int64_t n_cols, int64_t batch_cols, size_t extra_per_group, int64_t* out) {
    const int64_t G_all = (n_cols + 31) / 32;
    if (batch_cols <= 0) batch_cols = c->tune_batch_cols;
    if (batch_cols > 0) {
        *out = std::min(G_all, (batch_cols + 31) / 32);
        return PMFS_OK;
    }
    int64_t gmax = 0;
    int rc = mem_limited_groups(c, algo, G_all, extra_per_group, &gmax);
    if (rc) return rc;
    // (a batch is ordered through 32-bit cell prefix sums and 2^30 cells at most: bound the cell matrix as well)
    const int64_t max_regions = std::max<int64_t>(1, ((int64_t)1 << 30) / std::max(1, c->plan.n_nodes));
    gmax = std::min(gmax, max_regions * 32);
    if (gmax >= G_all) {
        *out = G_all;
        return PMFS_OK;
    }
    *out = balanced_batch_groups(G_all, gmax);
    return PMFS_OK;
}

int pmfs_run_batched(pmfs_ctx* c, int algo, const uint8_t* d_msa, int64_t row_stride, int32_t n_rows, int64_t n_cols,
                     const uint8_t* d_cons, const uint8_t* d_forced, int64_t batch_cols, pmfs_batch_fn consumer, void* user,
                     int64_t* n_total, int32_t* n_batches) {
    if (!c || !d_msa || !d_cons || n_cols <= 0 || batch_cols < 0) {
        g_error = "pmfs_run_batched: bad argument";
        return PMFS_ERR_ARG;
    }
    if (c->parts.n_parts > 0) {
        g_error = "pmfs_run_batched: not available while pmfs_set_shard node ranges are set (use pmfs_comm_run)";
        return PMFS_ERR_STATE;
    }
    PMFS_CUDA(cudaSetDevice(c->device));
    const bool fixed = batch_cols > 0 || c->tune_batch_cols > 0;
    int64_t gb = 0;
    int rc = batch_groups_for(c, algo, n_cols, batch_cols, 0, &gb);
    if (rc) return rc;
    const int64_t base_off = c->shard_total_cols > 0 ? c->shard_pos_offset : 0, base_total = c->shard_total_cols;
    int64_t total = 0;
    int32_t batches = 0;
    pmfs_timing sum = {};
    for (int64_t b0 = 0; b0 < n_cols;) {
        const int64_t nb = std::min(n_cols - b0, gb * 32);
        // a batch is a column shard that keeps its output: global positions, ordered by (node,pos) on its own
        c->shard_pos_offset = base_off + b0;
        c->shard_total_cols = base_total > 0 ? base_total : n_cols;
        pmfs_mut* d_out = nullptr;
        int64_t n = 0;
        c->timing.h2d_ms = 0;
        rc = run_device_impl(c, algo, d_msa + b0 / 2, row_stride, n_rows, nb, d_cons + b0, d_forced ? d_forced + b0 : nullptr, &d_out, &n);
        c->shard_pos_offset = base_off;
        c->shard_total_cols = base_total;
        if (rc == PMFS_ERR_ARG && c->too_many_records && !fixed && gb > 32) {
            gb = std::max<int64_t>(32, gb / 2 / 32 * 32);  // automatic batch size: halve it and run this batch again
            continue;
        }
        if (rc) return rc;
        const pmfs_timing& t = c->timing;
        sum.forward_ms += t.forward_ms; sum.backward_ms += t.backward_ms; sum.sort_ms += t.sort_ms; sum.total_ms += t.total_ms;
        sum.wall_ms += t.wall_ms; sum.n_launches += t.n_launches; sum.forward_bytes += t.forward_bytes;
        sum.backward_bytes += t.backward_bytes;
        if (consumer) {
            const int crc = consumer(user, base_off + b0, nb, d_out, n);
            if (crc != 0) {
                g_error = "pmfs_run_batched: the consumer returned " + std::to_string(crc);
                return PMFS_ERR_STATE;
            }
        }
        total += n;
        batches++;
        b0 += nb;
    }
    c->timing = sum;
    if (n_total) *n_total = total;
    if (n_batches) *n_batches = batches;
    return PMFS_OK;
}

int pmfs_set_shard(pmfs_ctx* c, int64_t pos_offset, int64_t total_cols, int32_t n_parts, const int32_t* node_bounds) {
    if (!c || n_parts < 0 || n_parts > 16 || pos_offset < 0 || total_cols < 0 || (n_parts > 0 && (!node_bounds || total_cols <= 0)) ||
        (total_cols == 0 && pos_offset != 0)) {
        g_error = "pmfs_set_shard: bad argument (at most 16 parts)";
        return PMFS_ERR_ARG;
    }
    c->parts.n_parts = n_parts;
    for (int r = 0; r <= n_parts && n_parts > 0; r++) {
        if (r > 0 && node_bounds[r] < node_bounds[r - 1]) {
            g_error = "pmfs_set_shard: node_bounds must not decrease";
            return PMFS_ERR_ARG;
        }
        c->parts.bound[r] = node_bounds[r];
    }
    c->shard_pos_offset = pos_offset;
    c->shard_total_cols = total_cols;
    return PMFS_OK;
}

int pmfs_get_parts(pmfs_ctx* c, const void** d_records, int64_t* offsets, int64_t* counts, int* column_ordered) {
    if (!c || !d_records || !offsets || !counts || !column_ordered || c->parts.n_parts <= 0) {
        g_error = "pmfs_get_parts: context is not in shard mode";
        return PMFS_ERR_STATE;
    }
    *d_records = c->parts_keys;
    *column_ordered = c->parts_by_columns ? 1 : 0;
    for (int r = 0; r < c->parts.n_parts; r++) {
        offsets[r] = c->part_offsets[r];
        counts[r] = c->part_counts[r];
    }
    return PMFS_OK;
}

int pmfs_sort_records(pmfs_ctx* c, void* d_records, int64_t n, int column_ordered, pmfs_mut** d_out) {
    if (!c || n < 0 || (n > 0 && !d_records) || !d_out || n > 0x7fffffffll) {
        g_error = "pmfs_sort_records: bad argument";
        return PMFS_ERR_ARG;
    }
    PMFS_CUDA(cudaSetDevice(c->device));
    const int64_t span = c->parts.n_parts > 0 ? c->shard_total_cols : 0;
    if (span <= 0) {
        g_error = "pmfs_sort_records: call pmfs_set_shard first (the record layout depends on the total column count)";
        return PMFS_ERR_STATE;
    }
    int pos_bits = 1, node_bits = 1;
    while ((1ll << pos_bits) < span) pos_bits++;
    while ((1ll << node_bits) < c->plan.n_nodes) node_bits++;
    int rc;
    if ((rc = c->d_out.ensure((size_t)std::max<int64_t>(1, n)))) return rc;
    if (n > 0) {
        if ((rc = c->d_rkeys_alt.ensure((size_t)n))) return rc;
        cub::DoubleBuffer<unsigned long long> kb((unsigned long long*)d_records, c->d_rkeys_alt.ptr);
        size_t tmp_bytes = 0;
        // blocks in source-rank order, each in column-region order with every node's records contiguous and
        // column-ordered inside a region: a STABLE sort on the node bits alone is enough (2 radix passes, not 4)
        const int lo = column_ordered ? kPayloadBits + pos_bits : kPayloadBits, hi = kPayloadBits + pos_bits + node_bits;
        PMFS_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, kb, (int)n, lo, hi, c->stream));
        if ((rc = c->d_sort_tmp2.ensure(tmp_bytes))) return rc;
        PMFS_CUDA(cub::DeviceRadixSort::SortKeys(c->d_sort_tmp2.ptr, tmp_bytes, kb, (int)n, lo, hi, c->stream));
        assemble_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(kb.Current(), n, pos_bits, c->d_out.ptr);
        PMFS_CUDA(cudaGetLastError());
    }
    PMFS_CUDA(cudaStreamSynchronize(c->stream));
    c->last_n_out = n;
    *d_out = c->d_out.ptr;
    return PMFS_OK;
}

// Greedy <=6-run packing of the last result ON THE DEVICE, then download of the packed NucMut lists.
int pmfs_pack_last(pmfs_ctx* c, const pmfs_host_nucmut_t** out, int64_t* n_out, const int64_t** node_off) {
    if (!c || !out || !n_out || !node_off) {
        g_error = "pmfs_pack_last: null argument";
        return PMFS_ERR_ARG;
    }
    PMFS_CUDA(cudaSetDevice(c->device));
    const int64_t n = c->last_n_out;
    const int32_t V = c->plan.n_nodes;
    int rc;
    if ((rc = c->d_pk_off.ensure((size_t)V + 1))) return rc;
    if (c->h_pk_off_cap < (size_t)V + 1) {
        if (c->h_pk_off) cudaFreeHost(c->h_pk_off);
        c->h_pk_off = nullptr;
        PMFS_CUDA(cudaMallocHost(&c->h_pk_off, ((size_t)V + 1) * sizeof(int64_t)));
        c->h_pk_off_cap = (size_t)V + 1;
    }
    int64_t n_packed = 0;
    if (n > 0) {
        if (n > 0x7fffffffll) {
            g_error = "pmfs_pack_last: too many records";
            return PMFS_ERR_ARG;
        }
        if ((rc = c->d_pk_run.ensure((size_t)n))) return rc;
        if ((rc = c->d_pk_idx.ensure((size_t)n + 1))) return rc;
        const unsigned nb = (unsigned)((n + 255) / 256);
        pack_runstart_kernel<<<nb, 256, 0, c->stream>>>(c->d_out.ptr, n, c->d_pk_run.ptr);
        size_t t1 = 0, t2 = 0;
        PMFS_CUDA(cub::DeviceScan::InclusiveScan(nullptr, t1, c->d_pk_run.ptr, c->d_pk_run.ptr, cub::Max(), (int)n, c->stream));
        PMFS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, t2, c->d_pk_idx.ptr, c->d_pk_idx.ptr, (int)n + 1, c->stream));
        if ((rc = c->d_pk_tmp.ensure(std::max(t1, t2)))) return rc;
        PMFS_CUDA(cub::DeviceScan::InclusiveScan(c->d_pk_tmp.ptr, t1, c->d_pk_run.ptr, c->d_pk_run.ptr, cub::Max(), (int)n, c->stream));
        pack_isstart_kernel<<<nb, 256, 0, c->stream>>>(c->d_pk_run.ptr, n, c->d_pk_idx.ptr);
        PMFS_CUDA(cudaMemsetAsync(c->d_pk_idx.ptr + n, 0, sizeof(int32_t), c->stream));
        PMFS_CUDA(cub::DeviceScan::ExclusiveSum(c->d_pk_tmp.ptr, t2, c->d_pk_idx.ptr, c->d_pk_idx.ptr, (int)n + 1, c->stream));
        int32_t total = 0;
        PMFS_CUDA(cudaMemcpyAsync(&total, c->d_pk_idx.ptr + n, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
        PMFS_CUDA(cudaStreamSynchronize(c->stream));
        n_packed = total;
        if ((rc = c->d_pk_out.ensure((size_t)n_packed))) return rc;
        if ((rc = c->d_pk_node.ensure((size_t)n_packed))) return rc;
        pack_emit_kernel<<<nb, 256, 0, c->stream>>>(c->d_out.ptr, c->d_pk_run.ptr, c->d_pk_idx.ptr, n, c->d_pk_out.ptr, c->d_pk_node.ptr);
    }
    pack_offsets_kernel<<<(unsigned)((n_packed + 256) / 256), 256, 0, c->stream>>>(c->d_pk_node.ptr, n_packed, V, c->d_pk_off.ptr);
    PMFS_CUDA(cudaGetLastError());
    if (c->h_pk_out_cap < (size_t)std::max<int64_t>(1, n_packed)) {
        if (c->h_pk_out) cudaFreeHost(c->h_pk_out);
        c->h_pk_out = nullptr;
        const size_t want = (size_t)n_packed + (size_t)n_packed / 4 + 1024;
        PMFS_CUDA(cudaMallocHost(&c->h_pk_out, want * sizeof(PackedNucMut)));
        c->h_pk_out_cap = want;
    }
    if (n_packed > 0)
        PMFS_CUDA(cudaMemcpyAsync(c->h_pk_out, c->d_pk_out.ptr, (size_t)n_packed * sizeof(PackedNucMut), cudaMemcpyDeviceToHost, c->stream));
    PMFS_CUDA(cudaMemcpyAsync(c->h_pk_off, c->d_pk_off.ptr, ((size_t)V + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, c->stream));
    PMFS_CUDA(cudaStreamSynchronize(c->stream));
    *out = (const pmfs_host_nucmut_t*)c->h_pk_out;
    *n_out = n_packed;
    *node_off = c->h_pk_off;
    return PMFS_OK;
}

int pmfs_download(pmfs_ctx* c, pmfs_mut* host_out, int64_t n) {
    if (!c || (!host_out && n > 0) || n < 0 || n > c->last_n_out) {
        g_error = "pmfs_download: bad argument";
        return PMFS_ERR_ARG;
    }
    if (n == 0) return PMFS_OK;
    PMFS_CUDA(cudaSetDevice(c->device));
    PMFS_CUDA(cudaMemcpyAsync(host_out, c->d_out.ptr, (size_t)n * sizeof(pmfs_mut), cudaMemcpyDeviceToHost, c->stream));
    PMFS_CUDA(cudaStreamSynchronize(c->stream));
    return PMFS_OK;
}

// pmfs_run when the whole alignment fits on the device next to (tiles of) its set arrays: one upload, one pass.
static int run_host_whole(pmfs_ctx* c, int algo, const uint8_t* msa, int64_t row_stride, int32_t n_rows, int64_t n_cols,
                          const uint8_t* cons, const uint8_t* forced, pmfs_mut** out, int64_t* n_out) {
    const int64_t dstride = ((n_cols + 31) / 32) * 16;
    int rc;
    if ((rc = c->d_msa.ensure((size_t)std::max<int64_t>(1, n_rows) * dstride))) return rc;
    if ((rc = c->d_cons.ensure((size_t)n_cols))) return rc;
    if (forced && (rc = c->d_forced.ensure((size_t)n_cols))) return rc;
    const int64_t used = (n_cols + 1) / 2;
    // Wide alignments go up in two column slabs on a copy stream and the passes of the first slab (all column tiles but
    // the last) run while the second is still on the bus: only the last tile's passes are left after the upload.
    constexpr int64_t kSlabGroups = 256;  // slab boundary granularity (a multiple of every tile shape and of an output region)
    const int64_t G_all = (n_cols + 31) / 32, tiles = (G_all + kSlabGroups - 1) / kSlabGroups;
    const bool slabs = n_rows > 0 && tiles >= 2 && (size_t)n_rows * (size_t)used >= ((size_t)32 << 20);
    pmfs_mut* d_out = nullptr;
    if (slabs) {
        const int64_t split = (tiles - 1) * kSlabGroups;  // column groups in slab 0
        const size_t w0 = (size_t)split * 16, w1 = (size_t)(row_stride >= dstride ? dstride : used) - w0;
        cudaStream_t cs = c->copy_stream;
        PMFS_CUDA(cudaEventRecord(c->ev[6], c->stream));  // whatever the caller's stream still does with these buffers
        PMFS_CUDA(cudaStreamWaitEvent(cs, c->ev[6], 0));
        PMFS_CUDA(cudaEventRecord(c->ev[5], cs));
        if (dstride > used && row_stride < dstride) PMFS_CUDA(cudaMemsetAsync(c->d_msa.ptr, 0, (size_t)n_rows * dstride, cs));
        PMFS_CUDA(cudaMemcpyAsync(c->d_cons.ptr, cons, (size_t)n_cols, cudaMemcpyHostToDevice, cs));
        if (forced) PMFS_CUDA(cudaMemcpyAsync(c->d_forced.ptr, forced, (size_t)n_cols, cudaMemcpyHostToDevice, cs));
        PMFS_CUDA(cudaMemcpy2DAsync(c->d_msa.ptr, (size_t)dstride, msa, (size_t)row_stride, w0, (size_t)n_rows, cudaMemcpyHostToDevice, cs));
        PMFS_CUDA(cudaEventRecord(c->ev_slab[0], cs));
        PMFS_CUDA(cudaMemcpy2DAsync(c->d_msa.ptr + w0, (size_t)dstride, msa + w0, (size_t)row_stride, w1, (size_t)n_rows,
                                    cudaMemcpyHostToDevice, cs));
        PMFS_CUDA(cudaEventRecord(c->ev_slab[1], cs));
        c->slab_split = split;
        rc = run_device_impl(c, algo, c->d_msa.ptr, dstride, n_rows, n_cols, c->d_cons.ptr, forced ? c->d_forced.ptr : nullptr,
                             &d_out, n_out);
        c->slab_split = 0;
        if (rc) {
            cudaStreamSynchronize(cs);
            return rc;
        }
        cudaEventElapsedTime(&c->timing.h2d_ms, c->ev[5], c->ev_slab[1]);  // (overlaps forward_ms / backward_ms)
        c->timing.total_ms = c->timing.wall_ms;
    } else {
        PMFS_CUDA(cudaEventRecord(c->ev[5], c->stream));
        if (dstride > used && row_stride != dstride) PMFS_CUDA(cudaMemsetAsync(c->d_msa.ptr, 0, (size_t)std::max<int64_t>(1, n_rows) * dstride, c->stream));
        if (n_rows > 0 && row_stride == dstride)
            PMFS_CUDA(cudaMemcpyAsync(c->d_msa.ptr, msa, (size_t)n_rows * dstride, cudaMemcpyHostToDevice, c->stream));
        else if (n_rows > 0)
            PMFS_CUDA(cudaMemcpy2DAsync(c->d_msa.ptr, (size_t)dstride, msa, (size_t)row_stride, (size_t)used, (size_t)n_rows,
                                        cudaMemcpyHostToDevice, c->stream));
        PMFS_CUDA(cudaMemcpyAsync(c->d_cons.ptr, cons, (size_t)n_cols, cudaMemcpyHostToDevice, c->stream));
        if (forced) PMFS_CUDA(cudaMemcpyAsync(c->d_forced.ptr, forced, (size_t)n_cols, cudaMemcpyHostToDevice, c->stream));
        PMFS_CUDA(cudaEventRecord(c->ev[6], c->stream));
        PMFS_CUDA(cudaEventSynchronize(c->ev[6]));
        cudaEventElapsedTime(&c->timing.h2d_ms, c->ev[5], c->ev[6]);
        // an odd final column shares its byte with a padding nibble that the caller may have left non-zero: it is masked
        // out of the emission by the real-column mask, and it cannot influence other columns.
        rc = run_device_impl(c, algo, c->d_msa.ptr, dstride, n_rows, n_cols, c->d_cons.ptr, forced ? c->d_forced.ptr : nullptr,
                             &d_out, n_out);
    }
    if (rc) return rc;
    // one pinned result buffer per context, grown on demand (pinning memory costs far more than the copy itself)
    if ((size_t)*n_out > c->h_out_cap || !c->h_out) {
        if (c->h_out) cudaFreeHost(c->h_out);
        c->h_out = nullptr;
        c->h_out_cap = 0;
        const size_t want = std::max<size_t>(1024, (size_t)*n_out + (size_t)*n_out / 4);
        PMFS_CUDA(cudaMallocHost(&c->h_out, want * sizeof(pmfs_mut)));
        c->h_out_cap = want;
    }
    pmfs_mut* h = c->h_out;
    PMFS_CUDA(cudaEventRecord(c->ev[5], c->stream));
    if (*n_out > 0)
        PMFS_CUDA(cudaMemcpyAsync(h, d_out, (size_t)*n_out * sizeof(pmfs_mut), cudaMemcpyDeviceToHost, c->stream));
    PMFS_CUDA(cudaEventRecord(c->ev[6], c->stream));
    PMFS_CUDA(cudaEventSynchronize(c->ev[6]));
    cudaEventElapsedTime(&c->timing.d2h_ms, c->ev[5], c->ev[6]);
    c->timing.total_ms += c->timing.d2h_ms;
    *out = h;
    return PMFS_OK;
}

// pmfs_run when it does not: the alignment goes through the GPU in column batches, batch b+1 on the bus while batch b is
// in the passes (two device buffers), every batch's ordered records come back to the host and are stitched into one
// (node,pos)-ordered array at the end -- a node's list is the concatenation of its slices in batch order.  The
// reference's --low-mem-mode has the same shape (src/panman.cpp:1503-1514: read 20 000 columns, process, append).
static int run_host_streaming(pmfs_ctx* c, int algo, const uint8_t* msa, int64_t row_stride, int32_t n_rows, int64_t n_cols,
                              const uint8_t* cons, const uint8_t* forced, pmfs_mut** out, int64_t* n_out) {
    const int64_t G_all = (n_cols + 31) / 32, used = (n_cols + 1) / 2;
    int rc;
    if ((rc = c->d_cons.ensure((size_t)n_cols))) return rc;
    if (forced && (rc = c->d_forced.ensure((size_t)n_cols))) return rc;
    c->d_msa.release();  // sized per batch below; what it held counts as free memory
    int64_t gb = 0;
    if (c->tune_batch_cols > 0) {
        gb = std::min(G_all, (c->tune_batch_cols + 31) / 32);
    } else {
        // a column group costs its share of the set arrays plus two staged MSA segments of 16 B per row
        size_t free_b = 0, total_b = 0;
        PMFS_CUDA(cudaMemGetInfo(&free_b, &total_b));
        free_b += c->d_S.cap * sizeof(uint4) + c->d_D.cap * sizeof(uint4);
        const int64_t per_group = (int64_t)c->plan.n_set_rows * 64 * (algo == PMFS_SANKOFF ? 2 : 1) + 64 + 2 * 16 * (int64_t)n_rows;
        int64_t gmax = std::max<int64_t>(1, (int64_t)(free_b * 0.65) / per_group);
        const int64_t max_regions = std::max<int64_t>(1, ((int64_t)1 << 30) / std::max(1, c->plan.n_nodes));
        gmax = std::min(gmax, max_regions * 32);
        gb = balanced_batch_groups(G_all, gmax);
    }
    const int64_t bstride = gb * 16;  // device row stride of a staged batch
    if ((rc = c->d_msa.ensure((size_t)2 * (size_t)std::max<int32_t>(1, n_rows) * (size_t)bstride))) return rc;
    uint8_t* const buf[2] = {c->d_msa.ptr, c->d_msa.ptr + (size_t)std::max<int32_t>(1, n_rows) * (size_t)bstride};
    cudaStream_t cs = c->copy_stream;
    PMFS_CUDA(cudaEventRecord(c->ev[6], c->stream));  // whatever the caller's stream still does with these buffers
    PMFS_CUDA(cudaStreamWaitEvent(cs, c->ev[6], 0));
    PMFS_CUDA(cudaMemcpyAsync(c->d_cons.ptr, cons, (size_t)n_cols, cudaMemcpyHostToDevice, cs));
    if (forced) PMFS_CUDA(cudaMemcpyAsync(c->d_forced.ptr, forced, (size_t)n_cols, cudaMemcpyHostToDevice, cs));
    const int64_t n_batches = (G_all + gb - 1) / gb;
    auto upload = [&](int64_t b) -> int {
        const int64_t g0 = b * gb, w = std::min<int64_t>(bstride, used - g0 * 16);  // bytes of a host row in this batch
        if (n_rows == 0) return PMFS_OK;
        if (w < bstride) PMFS_CUDA(cudaMemsetAsync(buf[b & 1], 0, (size_t)n_rows * bstride, cs));
        PMFS_CUDA(cudaMemcpy2DAsync(buf[b & 1], (size_t)bstride, msa + g0 * 16, (size_t)row_stride, (size_t)w, (size_t)n_rows,
                                    cudaMemcpyHostToDevice, cs));
        return PMFS_OK;
    };
    std::vector<pmfs_mut>& acc = c->h_acc;
    std::vector<int64_t> batch_end;
    acc.clear();
    const int64_t base_off = c->shard_total_cols > 0 ? c->shard_pos_offset : 0, base_total = c->shard_total_cols;
    pmfs_timing sum = {};
    const auto t0 = std::chrono::steady_clock::now();
    if ((rc = upload(0))) return rc;
    PMFS_CUDA(cudaEventRecord(c->ev_slab[0], cs));
    for (int64_t b = 0; b < n_batches; b++) {
        if (b + 1 < n_batches) {  // (its buffer is free: batch b-1 has been through the passes and its records are on the host)
            if ((rc = upload(b + 1))) return rc;
            PMFS_CUDA(cudaEventRecord(c->ev_slab[(b + 1) & 1], cs));
        }
        PMFS_CUDA(cudaStreamWaitEvent(c->stream, c->ev_slab[b & 1], 0));
        const int64_t g0 = b * gb, nb = std::min(n_cols - g0 * 32, gb * 32);
        c->shard_pos_offset = base_off + g0 * 32;
        c->shard_total_cols = base_total > 0 ? base_total : n_cols;
        pmfs_mut* d_out = nullptr;
        int64_t n = 0;
        c->timing.h2d_ms = 0;
        rc = run_device_impl(c, algo, buf[b & 1], bstride, n_rows, nb, c->d_cons.ptr + g0 * 32, forced ? c->d_forced.ptr + g0 * 32 : nullptr,
                             &d_out, &n);
        c->shard_pos_offset = base_off;
        c->shard_total_cols = base_total;
        if (rc) {
            cudaStreamSynchronize(cs);
            return rc;
        }
        const size_t at = acc.size();
        acc.resize(at + (size_t)n);
        if (n > 0) PMFS_CUDA(cudaMemcpyAsync(acc.data() + at, d_out, (size_t)n * sizeof(pmfs_mut), cudaMemcpyDeviceToHost, c->stream));
        PMFS_CUDA(cudaStreamSynchronize(c->stream));
        batch_end.push_back((int64_t)acc.size());
        const pmfs_timing& t = c->timing;
        sum.forward_ms += t.forward_ms; sum.backward_ms += t.backward_ms; sum.sort_ms += t.sort_ms;
        sum.n_launches += t.n_launches; sum.forward_bytes += t.forward_bytes; sum.backward_bytes += t.backward_bytes;
    }
    // stitch by node: count per node, prefix, then the batches in order (each is sorted by (node,pos))
    const size_t V = (size_t)c->plan.n_nodes, total = acc.size();
    std::vector<pmfs_mut>& res = c->h_big;
    if (n_batches == 1) {
        res.swap(acc);
    } else {
        std::vector<int64_t> at(V + 1, 0);
        for (size_t i = 0; i < total; i++) at[(size_t)acc[i].node + 1]++;
        for (size_t v = 0; v < V; v++) at[v + 1] += at[v];
        res.resize(total);
        for (size_t i = 0; i < total; i++) res[(size_t)at[(size_t)acc[i].node]++] = acc[i];
        std::vector<pmfs_mut>().swap(acc);
    }
    sum.wall_ms = sum.total_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    c->timing = sum;
    c->last_n_out = 0;  // the device only holds the last batch: pmfs_download / pmfs_pack_last do not apply
    *out = res.data();
    *n_out = (int64_t)total;
    return PMFS_OK;
}

int pmfs_run(pmfs_ctx* c, int algo, const uint8_t* msa, int64_t row_stride, int32_t n_rows, int64_t n_cols,
             const uint8_t* cons, const uint8_t* forced, pmfs_mut** out, int64_t* n_out) {
    if (!c || !msa || !cons || !out || !n_out || n_cols <= 0 || n_rows < 0 || row_stride * 2 < n_cols ||
        (algo != PMFS_FITCH && algo != PMFS_SANKOFF)) {
        g_error = "pmfs_run: bad argument";
        return PMFS_ERR_ARG;
    }
    if ((int64_t)c->plan.row_to_node.size() > (int64_t)n_rows) {
        g_error = "pmfs_run: a leaf refers to an MSA row >= n_rows";
        return PMFS_ERR_ARG;
    }
    PMFS_CUDA(cudaSetDevice(c->device));
    bool streaming = c->tune_batch_cols > 0 && c->tune_batch_cols < n_cols;
    if (!streaming && c->parts.n_parts == 0) {
        // whole upload unless the nibbles alone would take more than half of what is free (what the context holds counts)
        const size_t need = (size_t)std::max<int32_t>(1, n_rows) * (size_t)(((n_cols + 31) / 32) * 16);
        if (need > c->d_msa.cap) {
            size_t free_b = 0, total_b = 0;
            PMFS_CUDA(cudaMemGetInfo(&free_b, &total_b));
            streaming = need > (free_b + c->d_msa.cap) / 2;
        }
    }
    if (!streaming) {
        const int rc = run_host_whole(c, algo, msa, row_stride, n_rows, n_cols, cons, forced, out, n_out);
        if (!(rc == PMFS_ERR_ARG && c->too_many_records && c->parts.n_parts == 0)) return rc;
    }
    return run_host_streaming(c, algo, msa, row_stride, n_rows, n_cols, cons, forced, out, n_out);
}

static uint8_t host_code_of_char(char ch) {  // src/panman.cpp:83-118
    switch (ch) {
        case 'A': return 1;  case 'C': return 2;  case 'M': return 3;  case 'G': return 4;
        case 'R': return 5;  case 'S': return 6;  case 'V': return 7;  case 'T': return 8;
        case 'W': return 9;  case 'Y': return 10; case 'H': return 11; case 'K': return 12;
        case 'D': return 13; case 'B': return 14; case 'N': return 15;
        default: return 0;
    }
}

static int run_bytes(pmfs_ctx* c, int algo, bool raw, const char* msa_chars, int64_t row_stride, int32_t n_rows, int64_t n_cols,
                     const char* consensus_chars, const char* forced_chars, pmfs_mut** out, int64_t* n_out);

int pmfs_run_chars(pmfs_ctx* c, int algo, const char* msa_chars, int64_t row_stride, int32_t n_rows, int64_t n_cols,
                   const char* consensus_chars, const char* forced_chars, pmfs_mut** out, int64_t* n_out) {
    return run_bytes(c, algo, false, msa_chars, row_stride, n_rows, n_cols, consensus_chars, forced_chars, out, n_out);
}

int pmfs_run_codes(pmfs_ctx* c, int algo, const uint8_t* codes, int64_t row_stride, int32_t n_rows, int64_t n_cols,
                   const uint8_t* consensus_codes, const uint8_t* forced_codes, pmfs_mut** out, int64_t* n_out) {
    return run_bytes(c, algo, true, (const char*)codes, row_stride, n_rows, n_cols, (const char*)consensus_codes,
                     (const char*)forced_codes, out, n_out);
}

static int run_bytes(pmfs_ctx* c, int algo, bool raw, const char* msa_chars, int64_t row_stride, int32_t n_rows, int64_t n_cols,
                     const char* consensus_chars, const char* forced_chars, pmfs_mut** out, int64_t* n_out) {
    if (!c || !msa_chars || !consensus_chars || !out || !n_out || n_cols <= 0 || n_rows < 0 || row_stride < n_cols) {
        g_error = "pmfs_run_chars / pmfs_run_codes: bad argument";
        return PMFS_ERR_ARG;
    }
    PMFS_CUDA(cudaSetDevice(c->device));
    const int64_t dstride = ((n_cols + 31) / 32) * 16;
    const int64_t cstride = ((n_cols + 15) / 16) * 16;  // device character rows, 16-byte aligned
    int rc;
    if ((rc = c->d_msa.ensure((size_t)std::max<int64_t>(1, n_rows) * dstride))) return rc;
    if ((rc = c->d_chars.ensure((size_t)std::max<int64_t>(1, n_rows) * cstride + 16))) return rc;
    if ((rc = c->d_cons.ensure((size_t)n_cols))) return rc;
    if (forced_chars && (rc = c->d_forced.ensure((size_t)n_cols))) return rc;
    std::vector<uint8_t> cons((size_t)n_cols), forced(forced_chars ? (size_t)n_cols : 0);
    for (int64_t i = 0; i < n_cols; i++) cons[i] = raw ? (uint8_t)consensus_chars[i] & 15 : host_code_of_char(consensus_chars[i]);
    for (size_t i = 0; i < forced.size(); i++) forced[i] = raw ? (uint8_t)forced_chars[i] & 15 : host_code_of_char(forced_chars[i]);
    PMFS_CUDA(cudaEventRecord(c->ev[5], c->stream));
    if (n_rows > 0) {
        PMFS_CUDA(cudaMemcpy2DAsync(c->d_chars.ptr, (size_t)cstride, msa_chars, (size_t)row_stride, (size_t)n_cols, (size_t)n_rows,
                                    cudaMemcpyHostToDevice, c->stream));
        const int64_t total = (dstride / 4) * n_rows;
        if (raw) pack_chars_kernel<true><<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(c->d_chars.ptr, cstride, n_rows, n_cols, c->d_msa.ptr, dstride);
        else pack_chars_kernel<false><<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(c->d_chars.ptr, cstride, n_rows, n_cols, c->d_msa.ptr, dstride);
        PMFS_CUDA(cudaGetLastError());
    }
    PMFS_CUDA(cudaMemcpyAsync(c->d_cons.ptr, cons.data(), (size_t)n_cols, cudaMemcpyHostToDevice, c->stream));
    if (forced_chars) PMFS_CUDA(cudaMemcpyAsync(c->d_forced.ptr, forced.data(), (size_t)n_cols, cudaMemcpyHostToDevice, c->stream));
    PMFS_CUDA(cudaEventRecord(c->ev[6], c->stream));
    PMFS_CUDA(cudaEventSynchronize(c->ev[6]));
    cudaEventElapsedTime(&c->timing.h2d_ms, c->ev[5], c->ev[6]);
    pmfs_mut* d_out = nullptr;
    rc = run_device_impl(c, algo, c->d_msa.ptr, dstride, n_rows, n_cols, c->d_cons.ptr, forced_chars ? c->d_forced.ptr : nullptr,
                         &d_out, n_out);
    if (rc) return rc;
    if ((size_t)*n_out > c->h_out_cap || !c->h_out) {
        if (c->h_out) cudaFreeHost(c->h_out);
        c->h_out = nullptr;
        c->h_out_cap = 0;
        const size_t want = std::max<size_t>(1024, (size_t)*n_out + (size_t)*n_out / 4);
        PMFS_CUDA(cudaMallocHost(&c->h_out, want * sizeof(pmfs_mut)));
        c->h_out_cap = want;
    }
    if (*n_out > 0)
        PMFS_CUDA(cudaMemcpyAsync(c->h_out, d_out, (size_t)*n_out * sizeof(pmfs_mut), cudaMemcpyDeviceToHost, c->stream));
    PMFS_CUDA(cudaStreamSynchronize(c->stream));
    *out = c->h_out;
    return PMFS_OK;
}

// Characters row by row, straight from wherever the caller keeps its sequences (the reference keeps a std::map of
// std::strings, src/panman.cpp:1271-1315): the rows are staged through two pinned buffers, so the host touches every
// character once (no contiguous second copy of the alignment), the upload of one chunk of rows overlaps the staging of
// the next, the device converts to packed codes chunk by chunk, and only 0.5 B per cell stays resident.
int pmfs_run_char_rows(pmfs_ctx* c, int algo, const char* const* rows, int32_t n_rows, int64_t n_cols, const char* consensus_chars,
                       const char* forced_chars, pmfs_mut** out, int64_t* n_out) {
    if (!c || (!rows && n_rows > 0) || !consensus_chars || !out || !n_out || n_cols <= 0 || n_rows < 0) {
        g_error = "pmfs_run_char_rows: bad argument";
        return PMFS_ERR_ARG;
    }
    PMFS_CUDA(cudaSetDevice(c->device));
    const int64_t dstride = ((n_cols + 31) / 32) * 16;
    const int64_t cstride = ((n_cols + 15) / 16) * 16;
    const int64_t chunk_rows = std::max<int64_t>(1, std::min<int64_t>(std::max<int32_t>(1, n_rows), ((int64_t)32 << 20) / cstride));
    int rc;
    if ((rc = c->d_msa.ensure((size_t)std::max<int32_t>(1, n_rows) * dstride))) return rc;
    if ((rc = c->d_chars.ensure((size_t)2 * chunk_rows * cstride + 16))) return rc;
    if ((rc = c->d_cons.ensure((size_t)n_cols))) return rc;
    if (forced_chars && (rc = c->d_forced.ensure((size_t)n_cols))) return rc;
    const size_t stage_bytes = (size_t)chunk_rows * cstride;
    if (c->h_stage_cap < 2 * stage_bytes) {
        if (c->h_stage) cudaFreeHost(c->h_stage);
        c->h_stage = nullptr;
        c->h_stage_cap = 0;
        PMFS_CUDA(cudaMallocHost(&c->h_stage, 2 * stage_bytes));
        c->h_stage_cap = 2 * stage_bytes;
    }
    std::vector<uint8_t> cons((size_t)n_cols), forced(forced_chars ? (size_t)n_cols : 0);
    for (int64_t i = 0; i < n_cols; i++) cons[i] = host_code_of_char(consensus_chars[i]);
    for (size_t i = 0; i < forced.size(); i++) forced[i] = host_code_of_char(forced_chars[i]);
    PMFS_CUDA(cudaEventRecord(c->ev[5], c->stream));
    for (int64_t r0 = 0, k = 0; r0 < n_rows; r0 += chunk_rows, k++) {
        const int64_t nr = std::min<int64_t>(chunk_rows, n_rows - r0);
        uint8_t* h = c->h_stage + (k & 1) * stage_bytes;
        uint8_t* d = c->d_chars.ptr + (k & 1) * stage_bytes;
        if (k >= 2) PMFS_CUDA(cudaEventSynchronize(c->ev_slab[k & 1]));  // the copy out of this staging buffer has finished
        for (int64_t r = 0; r < nr; r++) memcpy(h + r * cstride, rows[r0 + r], (size_t)n_cols);
        PMFS_CUDA(cudaMemcpyAsync(d, h, (size_t)nr * cstride, cudaMemcpyHostToDevice, c->stream));
        PMFS_CUDA(cudaEventRecord(c->ev_slab[k & 1], c->stream));
        const int64_t total = (dstride / 4) * nr;
        pack_chars_kernel<false><<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(d, cstride, (int32_t)nr, n_cols, c->d_msa.ptr + r0 * dstride, dstride);
        PMFS_CUDA(cudaGetLastError());
    }
    PMFS_CUDA(cudaMemcpyAsync(c->d_cons.ptr, cons.data(), (size_t)n_cols, cudaMemcpyHostToDevice, c->stream));
    if (forced_chars) PMFS_CUDA(cudaMemcpyAsync(c->d_forced.ptr, forced.data(), (size_t)n_cols, cudaMemcpyHostToDevice, c->stream));
    PMFS_CUDA(cudaEventRecord(c->ev[6], c->stream));
    PMFS_CUDA(cudaEventSynchronize(c->ev[6]));
    cudaEventElapsedTime(&c->timing.h2d_ms, c->ev[5], c->ev[6]);
    pmfs_mut* d_out = nullptr;
    rc = run_device_impl(c, algo, c->d_msa.ptr, dstride, n_rows, n_cols, c->d_cons.ptr, forced_chars ? c->d_forced.ptr : nullptr, &d_out, n_out);
    if (rc) return rc;
    if ((size_t)*n_out > c->h_out_cap || !c->h_out) {
        if (c->h_out) cudaFreeHost(c->h_out);
        c->h_out = nullptr;
        c->h_out_cap = 0;
        const size_t want = std::max<size_t>(1024, (size_t)*n_out + (size_t)*n_out / 4);
        PMFS_CUDA(cudaMallocHost(&c->h_out, want * sizeof(pmfs_mut)));
        c->h_out_cap = want;
    }
    if (*n_out > 0) PMFS_CUDA(cudaMemcpyAsync(c->h_out, d_out, (size_t)*n_out * sizeof(pmfs_mut), cudaMemcpyDeviceToHost, c->stream));
    PMFS_CUDA(cudaStreamSynchronize(c->stream));
    *out = c->h_out;
    return PMFS_OK;
}

void pmfs_result_free(pmfs_ctx* c, pmfs_mut* out) {
    // results live in the context's pinned buffer, which is reused by the next pmfs_run and released by pmfs_destroy;
    // releasing early only matters for very large results
    if (c && out && out == c->h_big.data()) {
        std::vector<pmfs_mut>().swap(c->h_big);
        return;
    }
    if (!c || !out || out != c->h_out) return;
    if (c->h_out_cap > ((size_t)64 << 20) / sizeof(pmfs_mut)) {
        cudaFreeHost(c->h_out);
        c->h_out = nullptr;
        c->h_out_cap = 0;
    }
}

int pmfs_synth_msa(pmfs_ctx* c, const pmfs_synth_spec* spec, int64_t col_begin, int64_t n_cols, uint8_t* d_msa,
                   int64_t row_stride, uint8_t* d_cons) {
    if (!c || !spec || !d_msa || n_cols <= 0 || col_begin < 0 || col_begin % 8 != 0 || row_stride % 4 != 0 || row_stride * 2 < n_cols) {
        g_error = "pmfs_synth_msa: bad argument";
        return PMFS_ERR_ARG;
    }
    PMFS_CUDA(cudaSetDevice(c->device));
    const int32_t n_rows = (int32_t)c->plan.row_to_node.size();
    if (n_rows == 0) {
        g_error = "pmfs_synth_msa: tree has no leaf with an MSA row";
        return PMFS_ERR_ARG;
    }
    if (!c->synth_ready) {
        // nodes grouped by depth; internal nodes get a slot in their level's state buffer
        const size_t V = c->parent.size();
        std::vector<int32_t> depth(V, 0);
        int32_t maxd = 0;
        for (size_t v = 1; v < V; v++) maxd = std::max(maxd, depth[v] = depth[c->parent[v]] + 1);
        std::vector<int64_t>& off = c->synth_lvl_off;
        off.assign((size_t)maxd + 2, 0);
        for (size_t v = 0; v < V; v++) off[depth[v] + 1]++;
        for (int32_t d = 0; d <= maxd; d++) off[d + 1] += off[d];
        std::vector<int32_t> node(V), pslot(V), mslot(V), slot_of(V, -1);
        std::vector<int64_t> fill(off.begin(), off.end() - 1), n_int((size_t)maxd + 1, 0);
        c->synth_max_slots = 1;
        for (size_t v = 0; v < V; v++) {  // ascending ids: parents get their slots before their children
            const int32_t d = depth[v];
            const int64_t k = fill[d]++;
            node[k] = (int32_t)v;
            pslot[k] = v == 0 ? 0 : slot_of[c->parent[v]];
            if (c->leaf_row[v] == -1) {
                slot_of[v] = (int32_t)n_int[d]++;
                mslot[k] = slot_of[v];
                c->synth_max_slots = std::max<int64_t>(c->synth_max_slots, n_int[d]);
            } else {
                mslot[k] = c->leaf_row[v] >= 0 ? -(c->leaf_row[v] + 1) : 0x7fffffff;
            }
        }
        int rc;
        if ((rc = c->d_synth_node.ensure(V))) return rc;
        if ((rc = c->d_synth_pslot.ensure(V))) return rc;
        if ((rc = c->d_synth_mslot.ensure(V))) return rc;
        PMFS_CUDA(cudaMemcpy(c->d_synth_node.ptr, node.data(), V * sizeof(int32_t), cudaMemcpyHostToDevice));
        PMFS_CUDA(cudaMemcpy(c->d_synth_pslot.ptr, pslot.data(), V * sizeof(int32_t), cudaMemcpyHostToDevice));
        PMFS_CUDA(cudaMemcpy(c->d_synth_mslot.ptr, mslot.data(), V * sizeof(int32_t), cudaMemcpyHostToDevice));
        c->synth_ready = true;
    }
    auto thr = [](float p) -> uint32_t {
        double x = (double)p * 4294967296.0;
        return x >= 4294967295.0 ? 0xffffffffu : (x <= 0 ? 0u : (uint32_t)x);
    };
    SynthParams s;
    s.seed = spec->seed;
    s.t_sub = thr(spec->p_sub); s.t_del = thr(spec->p_del); s.t_ins = thr(spec->p_ins);
    s.t_nblock = thr(spec->p_n_block); s.t_amb = thr(spec->p_amb); s.t_acc = thr(spec->accessory_frac);
    s.words = (n_cols + 7) / 8;
    s.n_rows = n_rows; s.col_begin = col_begin; s.n_cols = n_cols; s.msa = d_msa; s.row_stride = row_stride;
    {
        int rc;
        if ((rc = c->d_synth_state[0].ensure((size_t)c->synth_max_slots * s.words))) return rc;
        if ((rc = c->d_synth_state[1].ensure((size_t)c->synth_max_slots * s.words))) return rc;
    }
    // (only this column window of the rows: the caller may be filling a wider alignment window by window)
    PMFS_CUDA(cudaMemset2DAsync(d_msa, (size_t)row_stride, 0, (size_t)s.words * 4, (size_t)n_rows, c->stream));
    for (size_t d = 0; d + 1 < c->synth_lvl_off.size(); d++) {
        const int64_t b0 = c->synth_lvl_off[d], e0 = c->synth_lvl_off[d + 1];
        s.node = c->d_synth_node.ptr + b0;
        s.parent_slot = c->d_synth_pslot.ptr + b0;
        s.my_slot = c->d_synth_mslot.ptr + b0;
        s.n_level = (int32_t)(e0 - b0);
        s.is_root = d == 0;
        s.prev = c->d_synth_state[(d + 1) & 1].ptr;
        s.cur = c->d_synth_state[d & 1].ptr;
        const int64_t total = s.words * s.n_level;
        synth_level_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(s);
    }
    if (d_cons) synth_consensus_kernel<<<(unsigned)((n_cols + 255) / 256), 256, 0, c->stream>>>(d_msa, row_stride, n_rows, n_cols, d_cons);
    PMFS_CUDA(cudaGetLastError());
    PMFS_CUDA(cudaStreamSynchronize(c->stream));
    if (c->d_synth_state[0].cap * sizeof(uint32_t) > ((size_t)256 << 20)) {  // big trees: do not sit on gigabytes of scratch
        c->d_synth_state[0].release();
        c->d_synth_state[1].release();
    }
    return PMFS_OK;
}

}  // extern "C"

// =====================================================================================================================
// Allocation-free Fitch run.
//
// Who this is for: code that must not allocate device memory at the moment it needs an answer -- a service loop
// with a latency budget (cudaMalloc/cudaFree synchronise the whole device and can take milliseconds), a host that has
// handed all of its HBM to another allocator and has nothing left to ask the driver for, or a region of the program in
// which the CUDA allocator may simply not be called.  pmfs_run_device grows and reuses buffers on demand and may
// therefore call cudaMalloc; this entry point never does.  Everything it needs beyond the caller's inputs and outputs
// (the per-node set array the two walk kernels hand to each other, the per-chunk parent states, the bit-sliced
// consensus, the unsorted record pairs and the radix sort's double buffers) is carved out of ONE array with static
// storage duration in device memory.  The price: the array has a fixed size, it is resident on every GPU that loads
// this library until the process exits, and all calls share it.
// =====================================================================================================================
namespace pmfs {

constexpr size_t kStaticScratchBytes = (size_t)32 << 20;  // 32 MiB per GPU that loads the library

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// In host code `g_static_scratch` is only a handle for the runtime, not a device address: handing it to a kernel would
// fault (or quietly touch host memory on systems where the GPU can reach it).  The real address differs per device and
// is looked up once per device with cudaGetSymbolAddress, with that device current.
constexpr int kMaxStaticDevices = 64;
static std::mutex g_static_mu;
static unsigned char* g_static_dev_ptr[kMaxStaticDevices] = {};

static int static_scratch_on(int device, unsigned char** out) {
    if (device < 0 || device >= kMaxStaticDevices) {
        g_error = "pmfs_run_fitch_static: device index out of range";
        return PMFS_ERR_ARG;
    }
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_dev_ptr[device]) {
        PMFS_CUDA(cudaSetDevice(device));
        void* ptr = nullptr;
        PMFS_CUDA(cudaGetSymbolAddress(&ptr, g_static_scratch));
        if (!ptr) {
            g_error = "pmfs_run_fitch_static: cudaGetSymbolAddress returned a null address";
            return PMFS_ERR_CUDA;
        }
        g_static_dev_ptr[device] = (unsigned char*)ptr;
    }
    *out = g_static_dev_ptr[device];
    return PMFS_OK;
}

struct Carver {  // hands out 256-byte aligned pieces of the static array
    unsigned char* base;
    size_t used = 0;
    template <class T>
    T* take(size_t n) {
        T* p = (T*)(base + used);
        used += (n * sizeof(T) + 255) / 256 * 256;
        return p;
    }
};

}  // namespace pmfs

extern "C" int pmfs_run_fitch_static(pmfs_ctx* c, const uint8_t* d_msa, int64_t row_stride, int32_t n_rows, int64_t n_cols,
                                     const uint8_t* d_cons, const uint8_t* d_forced, pmfs_mut* d_out, int64_t out_capacity,
                                     int64_t* n_out) {
    if (!c || !d_msa || !d_cons || !d_out || !n_out || n_cols <= 0 || out_capacity <= 0) {
        g_error = "pmfs_run_fitch_static: bad argument";
        return PMFS_ERR_ARG;
    }
    if (row_stride % 16 != 0 || ((uintptr_t)d_msa & 15) != 0 || row_stride * 2 < n_cols) {
        g_error = "pmfs_run_fitch_static: MSA rows must be 16-byte aligned with a stride that is a multiple of 16 and covers n_cols";
        return PMFS_ERR_ARG;
    }
    if ((int64_t)c->plan.row_to_node.size() > (int64_t)n_rows) {
        g_error = "pmfs_run_fitch_static: a leaf refers to an MSA row >= n_rows";
        return PMFS_ERR_ARG;
    }
    if (c->parts.n_parts > 0 || c->shard_total_cols > 0) {
        g_error = "pmfs_run_fitch_static: not available on a context in column-shard mode (pmfs_set_shard)";
        return PMFS_ERR_STATE;
    }
    *n_out = 0;
    c->last_n_out = 0;  // the result goes to the caller's buffer: pmfs_download / pmfs_pack_last have nothing to work on
    unsigned char* scratch = nullptr;
    int rc = static_scratch_on(c->device, &scratch);  // first call on a device: resolves the array's address there
    if (rc) return rc;
    PMFS_CUDA(cudaSetDevice(c->device));

    if (c->plan.n_parts > 0 && c->plan.part_rows != kPartRowsFitch) {
        // (re-planning would allocate; this entry point never does)
        g_error = "pmfs_run_fitch_static: the tree has split polytomies and the context's plan was last built for Sankoff: "
                  "run pmfs_run_device(PMFS_FITCH) once first";
        return PMFS_ERR_STATE;
    }
    const Plan& P = c->plan;
    const int64_t G = (n_cols + 31) / 32;
    const size_t I = (size_t)P.n_set_rows, NC = P.chunks.size(), cap = (size_t)out_capacity;
    int node_bits = 1, pos_bits = 1;
    while ((1ll << node_bits) < P.n_nodes) node_bits++;
    while ((1ll << pos_bits) < n_cols) pos_bits++;
    size_t sort_tmp = 0;
    if (out_capacity > 0x7fffffffll) {
        g_error = "pmfs_run_fitch_static: out_capacity too large";
        return PMFS_ERR_ARG;
    }
    if (pos_bits + node_bits + kPayloadBits > 64) {
        g_error = "pmfs_run_fitch_static: log2(nodes) + log2(columns) must not exceed 54";
        return PMFS_ERR_ARG;
    }
    {
        cub::DoubleBuffer<unsigned long long> kb(nullptr, nullptr);
        PMFS_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, sort_tmp, kb, (int)cap, kPayloadBits, kPayloadBits + pos_bits + node_bits, c->stream));
    }
    Carver cv{scratch};
    uint4* S = cv.take<uint4>(I * 4 * (size_t)G);
    uint4* xcode = cv.take<uint4>(NC * (size_t)G);
    uint32_t* xvalid = cv.take<uint32_t>(NC * (size_t)G);
    uint4* cons_planes = cv.take<uint4>((size_t)G);
    uint4* forced_planes = cv.take<uint4>((size_t)G);
    unsigned long long* count = cv.take<unsigned long long>(1);
    unsigned int* err = cv.take<unsigned int>(1);
    unsigned long long* keys = cv.take<unsigned long long>(cap);
    unsigned long long* keys_alt = cv.take<unsigned long long>(cap);
    uint8_t* tmp = cv.take<uint8_t>(sort_tmp);
    c->ks = choose_shape(c, G);
    int* sync = cv.take<int>(sync_ints(P, G, c->ks->shape.groups()));
    if (cv.used > kStaticScratchBytes) {
        g_error = "pmfs_run_fitch_static: this problem needs " + std::to_string(cv.used) + " bytes of scratch (" +
                  std::to_string(I) + " internal nodes x " + std::to_string(G) + " column groups x 64 B, plus 16 B per record "
                  "of out_capacity); the static maximum is " + std::to_string(kStaticScratchBytes) +
                  " bytes.  Use pmfs_run_device, which sizes its buffers to the problem.";
        return PMFS_ERR_ARG;
    }

    // the array keeps whatever the previous call left in it: the record counter and the error flag must start at zero;
    // every set / parent-state entry that is read is written earlier in the same call, so those need no clearing
    PMFS_CUDA(cudaMemsetAsync(count, 0, sizeof(unsigned long long), c->stream));
    PMFS_CUDA(cudaMemsetAsync(err, 0, sizeof(unsigned int), c->stream));

    WalkParams wp = {};
    wp.G = G;
    wp.col_base = 0;
    wp.tile_cols = n_cols;
    wp.msa = (const uint4*)d_msa;
    wp.msa_stride16 = row_stride / 16;
    wp.S = S;
    wp.D = nullptr;
    wp.cons_planes = cons_planes;
    wp.forced_planes = d_forced ? forced_planes : nullptr;
    wp.xcode = xcode;
    wp.xvalid = xvalid;
    wp.int_to_node = c->d_int_to_node.ptr;
    wp.row_to_node = c->d_row_to_node.ptr;
    wp.out_keys = keys;
    wp.n_owners = 1;
    wp.out_count = count;
    wp.out_cap = cap;
    wp.err_flag = err;
    wp.pos_bits = pos_bits;
    wp.ready = sync;  // claim and dependency counters: zeroed by launch_levels before each pass
    const unsigned pg = (unsigned)((G + 127) / 128);
    prep_planes_kernel<<<pg, 128, 0, c->stream>>>(d_cons, n_cols, G, cons_planes);
    if (d_forced) prep_planes_kernel<<<pg, 128, 0, c->stream>>>(d_forced, n_cols, G, forced_planes);
    int64_t launches = 0;
    if ((rc = launch_levels<kAlgoFitch>(c, wp, true, &launches))) return rc;
    if ((rc = launch_levels<kAlgoFitch>(c, wp, false, &launches))) return rc;
    unsigned long long n = 0;
    PMFS_CUDA(cudaMemcpyAsync(&n, count, sizeof(n), cudaMemcpyDeviceToHost, c->stream));
    PMFS_CUDA(cudaStreamSynchronize(c->stream));
    *n_out = (int64_t)n;
    if (n > cap) {
        g_error = "pmfs_run_fitch_static: " + std::to_string(n) + " records do not fit in out_capacity " + std::to_string(cap);
        return PMFS_ERR_ARG;
    }
    if (n > 0) {
        cub::DoubleBuffer<unsigned long long> kb(keys, keys_alt);
        PMFS_CUDA(cub::DeviceRadixSort::SortKeys(tmp, sort_tmp, kb, (int)n, kPayloadBits, kPayloadBits + pos_bits + node_bits, c->stream));
        assemble_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(kb.Current(), (int64_t)n, pos_bits, d_out);
        PMFS_CUDA(cudaGetLastError());
    }
    PMFS_CUDA(cudaStreamSynchronize(c->stream));
    return PMFS_OK;
}
