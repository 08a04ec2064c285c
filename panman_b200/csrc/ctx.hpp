// What the translation units of libpmfs share: error text, the context, and the internal entry points of the pipeline
// (pmfs.cu) that the multi-GPU layer (comm.cu) builds on.
#pragma once

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/pmfs.h"
#include "plan.hpp"
#include "walk_params.hpp"

namespace pmfs {

extern thread_local std::string g_error;  // text of pmfs_last_error(), per host thread

#define PMFS_CUDA(call)                                                                             \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess) {                                                                    \
            g_error = std::string(#call) + ": " + cudaGetErrorString(e_);                           \
            return e_ == cudaErrorMemoryAllocation ? PMFS_ERR_NOMEM : PMFS_ERR_CUDA;                \
        }                                                                                           \
    } while (0)


struct PackedNucMut {  // pmfs_host_nucmut (include/pmfs_host.h)
    int32_t nucPosition, nucGapPosition, primaryBlockId, secondaryBlockId;
    uint32_t mutInfo, nucs;
};

// ------------------------------------------------------------------------------------------------ context
template <class T>
struct DevBuf {
    T* ptr = nullptr;
    size_t cap = 0;  // elements
    int ensure(size_t n) {
        if (n <= cap) return PMFS_OK;
        if (ptr) cudaFree(ptr);
        ptr = nullptr;
        cap = 0;
        cudaError_t e = cudaMalloc(&ptr, std::max<size_t>(n, 1) * sizeof(T));
        if (e != cudaSuccess) {
            g_error = std::string("cudaMalloc(") + std::to_string(n * sizeof(T)) + " B): " + cudaGetErrorString(e);
            cudaGetLastError();
            return PMFS_ERR_NOMEM;
        }
        cap = n;
        return PMFS_OK;
    }
    void release() {
        if (ptr) cudaFree(ptr);
        ptr = nullptr;
        cap = 0;
    }
};

}  // namespace pmfs

using namespace pmfs;  // (internal header: only the library's own translation units include it)

struct PartBounds {
    int32_t n_parts;
    int32_t bound[17];  // nodes [bound[r], bound[r+1]) belong to rank r
};

struct pmfs_ctx {
    int device = 0;
    Plan plan;
    bool has_poly = false;
    int32_t chunk_nodes_used = 0;
    bool chunk_nodes_auto = false;  // the plan in use was made for an automatic (not a caller-given) chunk size
    std::vector<int32_t> parent, child_off, child_idx, leaf_row;  // kept for re-planning and synth
    int32_t tune_chunk_nodes = 0;
    int32_t tune_split_arity = 128, tune_part_children = 63;  // split polytomies (PlanOptions)
    int64_t tune_tile_cols = 0;
    int tune_nw = 0, tune_ng = 0;          // pmfs_set_tile: forced tile shape (0 = automatic)
    const WalkKernelSet* ks = nullptr;     // tile shape of the current run
    int32_t tree_height = 0;               // internal nodes on the longest root-to-leaf path
    long long watchdog_ns = 10000000000ll; // device-side wait limit (pmfs_set_watchdog)
    cudaStream_t stream = nullptr;      // stream in use
    cudaStream_t own_stream = nullptr;  // created by pmfs_create
    cudaEvent_t ev[8] = {};
    cudaEvent_t ev_slab[2] = {};        // pmfs_run: upload slab 0 / 1 complete (recorded on copy_stream)
    cudaStream_t copy_stream = nullptr;
    int64_t slab_split = 0;             // > 0 while pmfs_run pipelines: first column group of the second slab
    int64_t cells_clean_n = 0;          // d_cell_count holds this many zeroed cells (0: unknown, clear before use)
    // plan on device
    DevBuf<uint32_t> d_fwd_prog, d_bwd_prog, d_fwd_stream, d_bwd_stream;
    DevBuf<int64_t> d_fwd_off, d_bwd_off, d_fwd_soff, d_bwd_soff;
    DevBuf<int32_t> d_fwd_slen, d_bwd_slen, d_chunk_parent, d_chunk_nchild;
    DevBuf<int4> d_fwd_items, d_bwd_items;
    DevBuf<int> d_sync;
    int n_sms = 0;
    struct Occ {
        const void* k;
        size_t smem;
        int per_sm;
    };
    std::vector<Occ> occ_cache;
    float host_ms = 0;  // wall time of the last run_device_impl call
    bool too_many_records = false;  // the last run failed only because one call may not produce that many records
    int64_t tune_batch_cols = 0;    // pmfs_set_batch_cols
    // column-shard mode (pmfs_set_shard)
    int64_t shard_pos_offset = 0, shard_total_cols = 0;
    PartBounds parts = {};
    int64_t part_counts[16] = {};
    unsigned long long* parts_keys = nullptr;
    int64_t part_offsets[16] = {};
    bool parts_by_columns = false;
    DevBuf<unsigned long long> d_rkeys_alt;
    DevBuf<uint8_t> d_sort_tmp2;
    DevBuf<int32_t> d_fwd_len, d_bwd_len, d_fwd_order, d_bwd_order, d_int_to_node, d_row_to_node;
    // per-run state
    DevBuf<uint4> d_S, d_D, d_cons_planes, d_forced_planes, d_xcode;
    DevBuf<uint32_t> d_xvalid;
    DevBuf<unsigned long long> d_keys, d_keys_alt;

    DevBuf<unsigned long long> d_count;
    DevBuf<uint8_t> d_sort_tmp, d_scan_tmp;
    DevBuf<uint32_t> d_cell_count, d_cell_start, d_cell_prefix;  // per (node, column region), see order_cells_kernel
    unsigned long long* h_stats = nullptr;  // pinned: {records, largest area, error flag} of the last run
    DevBuf<pmfs_mut> d_out;
    // gathered column shards (comm.cu): the ordered result stays 64-bit records, with per-node offsets and owner cuts
    bool keys_out = false;
    PartBounds xparts = {};                 // node ranges of the ranks
    DevBuf<unsigned long long> d_skeys;     // ordered records
    const unsigned long long* skeys_view = nullptr;  // where they are (d_skeys, or a sort buffer on the radix path)
    DevBuf<uint32_t> d_noff;                // [V+1] first record of every node
    DevBuf<unsigned long long> d_cuts;      // [n_parts+1] first record of every rank's node range
    unsigned long long* h_cuts = nullptr;   // pinned copy
    // host-path staging
    DevBuf<uint8_t> d_msa, d_cons, d_forced, d_chars;
    std::vector<unsigned long long> h_counts, h_offsets;
    // device-side packing (pmfs_pack_last)
    DevBuf<int32_t> d_pk_run, d_pk_idx, d_pk_node;
    DevBuf<PackedNucMut> d_pk_out;
    DevBuf<int64_t> d_pk_off;
    DevBuf<uint8_t> d_pk_tmp;
    PackedNucMut* h_pk_out = nullptr;  // pinned
    int64_t* h_pk_off = nullptr;       // pinned
    size_t h_pk_out_cap = 0, h_pk_off_cap = 0;
    pmfs_mut* h_out = nullptr;  // pinned
    size_t h_out_cap = 0;
    uint8_t* h_stage = nullptr;  // pinned: two staging buffers of pmfs_run_char_rows
    size_t h_stage_cap = 0;
    std::vector<pmfs_mut> h_acc, h_big;  // pmfs_run in column batches: per-batch records, and the stitched result
    int64_t last_n_out = 0;
    pmfs_timing timing = {};
    // synth
    DevBuf<int32_t> d_synth_node, d_synth_pslot, d_synth_mslot;
    DevBuf<uint32_t> d_synth_state[2];
    std::vector<int64_t> synth_lvl_off;
    int64_t synth_max_slots = 1;
    bool synth_ready = false;
};


// ---- internal entry points (pmfs.cu) ----------------------------------------------------------------------------------
namespace pmfs {

// The whole pass over one column window, device pointers in, records on the device out (what pmfs_run_device wraps).
// With c->keys_out set the ordered result is left as 64-bit records in c->d_skeys (and nothing is written to d_out).
int run_device_impl(pmfs_ctx* c, int algo, const uint8_t* d_msa, int64_t row_stride, int32_t n_rows, int64_t n_cols,
                    const uint8_t* d_cons, const uint8_t* d_forced, pmfs_mut** d_out, int64_t* n_out);
// Column groups whose set arrays fit in the device memory that is free now.
int mem_limited_groups(pmfs_ctx* c, int algo, int64_t G_all, size_t extra_per_group, int64_t* groups);
// Equal batches of at most gmax column groups, sized in multiples of a wide tile shape.
int64_t balanced_batch_groups(int64_t G_all, int64_t gmax);

}  // namespace pmfs
