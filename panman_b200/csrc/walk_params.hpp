// What the host side (pmfs.cu) and the walk kernels (walk_kernels.cuh, one translation unit per column-tile shape)
// share: the kernel parameter block and the table through which pmfs.cu reaches the kernels of one tile shape.
#pragma once

#include <cstddef>
#include <cstdint>

#include <vector_types.h>

namespace pmfs {

constexpr int kPayloadBits = 10;  // low bits of a 64-bit record: type 2, node state 4, parent state 4 (pmfs.cu)

struct WalkParams {
    const uint32_t* prog;       // fwd or bwd program words
    const int64_t* chunk_off;   // [NC] offset of each chunk's program
    const int32_t* chunk_len;   // [NC]
    const int32_t* order;       // chunk ids in execution order of this pass
    const int4* items;          // [2*NC] per chunk IN EXECUTION ORDER: {chunk, parent, plen, slen}, {prog_off/4, stream_off, need, dep chunk}
    const int32_t* chunk_parent;  // [NC] parent chunk, -1 for the chunk that holds the tree root
    const int32_t* chunk_nchild;  // [NC] number of child chunks
    int* ready;                 // [NC][ctas_per_chunk] dependency counters of this pass (zeroed before the launch)
    int* next_item;             // work-item claim counter of this pass (zeroed before the launch)
    int32_t n_items;
    const uint32_t* stream;     // load streams (plan.hpp)
    const int64_t* stream_off;  // [NC]
    const int32_t* stream_len;  // [NC]
    int32_t ctas_per_chunk;
    int32_t stack_depth;
    int32_t claim_lead;         // stream entries before the end of an item at which the producer claims the next one
    int32_t leaf_ring_log2;     // operand rings of a CTA: 2^n leaf-row units and 2^n set slots in flight (walk_kernels.cuh)
    int32_t set_ring_log2;
    const uint4* msa;           // first group of this column tile
    int64_t msa_stride16;       // uint4 per MSA row
    uint4* S;                   // [I][4][G]
    uint4* D;                   // Sankoff only
    int64_t G;                  // column groups in this tile
    int64_t tile_cols;          // real columns in this tile
    int64_t col_base;           // global column of the tile's first column
    const uint4* cons_planes;   // [G]
    const uint4* forced_planes; // [G] or null
    uint4* xcode;               // [NC][G] parent state per chunk root
    uint32_t* xvalid;           // [NC][G]
    const int32_t* int_to_node;
    const int32_t* row_to_node;
    unsigned long long* out_keys;   // [n_owners][out_cap] packed records (see kPayloadBits)
    unsigned long long* out_count;  // [n_owners]
    unsigned long long out_cap;     // capacity of one owner's region
    int32_t n_regions;              // > 0: one output region per 32 column groups (see emit); 0: regions per owner
    uint32_t bucket_mask;           // single-GPU runs: regions are further split by (node & bucket_mask), bucket-major
    uint32_t* cell_count;           // [n_nodes][n_regions] or null: records of (node, region) -- see order_cells_kernel
    uint32_t* cell_start;           // [n_nodes][n_regions]: where they start inside the region
    int64_t region_base;            // region index of this column tile's first 32 groups
    int32_t n_owners;               // 1, or the number of ranks in column-shard mode
    int32_t owner_bound[17];        // nodes [owner_bound[r], owner_bound[r+1]) are owned by rank r
    unsigned int* err_flag;
    int64_t pos_offset;         // global column of this context's first column (column shards); 0 otherwise
    int pos_bits;               // record = ((node << pos_bits | column) << kPayloadBits) | payload
    long long spin_limit_ns;    // watchdog of every device-side wait (0 = wait forever), see pmfs_set_watchdog
};

// Shape of a work item's column tile: NW walking warps, each thread walking NG column groups of 32 columns, plus one
// producer warp.  A tile is NW*32*NG column groups.  Wide tiles amortise the per-item and per-op bookkeeping; narrow
// tiles give more independent work items when the tree offers few parallel chunks (a ladder) and waste fewer lanes on
// narrow column ranges (column shards of a wide tree, memory-limited column batches).
struct TileShape {
    int nw, ng;
    int groups() const { return nw * 32 * ng; }
    int threads() const { return nw * 32 + 32; }
};

// The kernels of one tile shape (defined in walk_inst_*.cu, one translation unit each).
struct WalkKernelSet {
    TileShape shape;
    const void* fwd[2][2];  // [algo][poly]
    const void* bwd[2];     // [algo]
    size_t smem_fixed;                              // dynamic shared memory without rings and walk stack
    size_t leaf_unit_bytes, set_slot_bytes;         // one ring entry of each kind
    size_t fwd_smem_per_level, bwd_smem_per_level;  // walk stack bytes per saved state
    // -DPMFS_WAITPROF builds only (null otherwise): clear / print the wait-time counters and dump the item timelines
    void (*prof_reset)();
    void (*prof_dump)(bool fwd, unsigned grid, int per_sm, int n_items);
};

constexpr int kNumTileShapes = 6;
const WalkKernelSet* walk_kernel_set(int nw, int ng);  // nullptr if that shape is not built
const WalkKernelSet* walk_kernel_set_by_index(int i);  // 0 .. kNumTileShapes-1, widest first

}  // namespace pmfs
