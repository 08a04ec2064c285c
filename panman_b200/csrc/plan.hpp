// Host-side execution plan for one tree: the flat, device-friendly stand-in for the Node* graph the reference walks
// recursively (src/fitchSankoff.cpp:30-171, :359-531, :676-703 in /root/reference).
//
// The tree is cut into CHUNKS (connected pieces, each rooted at an internal node).  For every chunk the plan holds two
// straight-line "walk programs" -- one bottom-up (post-order), one top-down (pre-order) -- that every GPU thread
// executes in lock step on its own 32 alignment columns: the tree is the instruction stream, the columns are the lanes.
// Chunks whose inputs come from other chunks are ordered by levels (height for the bottom-up pass, depth for the
// top-down pass); chunks of one level are independent and run as one grid.
#pragma once

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/pmfs.h"

namespace pmfs {

// ---- walk-program encoding: one 32-bit word per op = opcode (low 3 bits) | argument << 3; some ops are followed by
// operand words.  A forward operand word names a child's state set: bit 31 set -> row `w & 0x7fffffff` of the
// internal-node set array, else MSA row `w`; kAbsentLeaf is a leaf with no MSA row (empty set).
constexpr uint32_t kOpBits = 3;
constexpr uint32_t kOperandSet = 0x80000000u;
constexpr uint32_t kOperandExt = 0x40000000u;  // with kOperandSet: the set was written by an EARLIER launch (root of a child
                                               // chunk), so it may be prefetched; without it the set belongs to this chunk
constexpr uint32_t kOperandIdMask = 0x3fffffffu;
constexpr uint32_t kAbsentLeaf = 0x3fffffffu;

// ---- load streams: next to each walk program, the sequence of operands the program will fetch from global memory,
// in consumption order, so that a producer warp can run them ahead of the walk through a shared-memory ring (TMA).
// Entry: bit 31 set -> the whole set row `e & 0x7fffffff` (4 quarter copies); else MSA row `e` (1 copy).
constexpr uint32_t kStreamSet = 0x80000000u;

enum FwdOp : uint32_t {
    F_PUSH = 0,       // push the running set T on the per-thread stack
    F_JOIN_POP = 1,   // arg = internal id v:  T = join(pop(), T);            store set[v]
    F_JOIN_IMM = 2,   // arg = v, 1 operand:   T = join(T, operand);          store set[v]
    F_JOIN_IMM2 = 3,  // arg = v, 2 operands:  T = join(operand0, operand1);  store set[v]
    F_POLY = 4,       // arg = v, next word = k, then k operands: any arity (1, >=3); children re-read from memory
    // Split polytomies (arity > PlanOptions::split_arity): the children are dealt out to PART chunks, each of which
    // folds its share into a partial result -- Fitch: the pair (AND, OR) of the children's sets; Sankoff: the bit-sliced
    // counters h[s] of its children (row b holds bit b of all 16 counters) -- kept in `part_rows` extra rows of the set
    // array; the node itself only combines the partial results.  fitchSankoff.cpp:39-55 (AND else OR over all
    // children) and :372-403 (sums over all children) are associative in exactly these partial forms.
    F_PART = 5,       // arg = first extra row, next word = k, then k operands: fold and store the partial result
    F_COMBINE = 6,    // arg = v, next word = m, then m operands (first extra row of each part): combine; store set[v]
};

enum BwdOp : uint32_t {
    B_ENTER = 0,       // arg = internal id v: resolve v against the parent state P, emit, then P := state(v)
    B_ENTER_ROOT = 1,  // arg = 0: the tree root (P holds the consensus)
    B_LEAF = 2,        // arg = MSA row: resolve + emit a present leaf against P
    B_XCHILD = 3,      // arg = chunk id: publish P for the child chunk rooted below this node
    B_PUSHP = 4,       // save P before descending into a non-last subtree
    B_POPP = 5,        // restore P
};

struct Chunk {
    int32_t root_node = 0;  // node id of the chunk root (always internal)
    int32_t root_int = 0;   // its internal id
    int64_t fwd_off = 0, bwd_off = 0;
    int32_t fwd_len = 0, bwd_len = 0;
    int64_t fwd_stream_off = 0, bwd_stream_off = 0;
    int32_t fwd_stream_len = 0, bwd_stream_len = 0;
    int32_t parent = -1;   // chunk that contains the parent of root_node; -1 for the chunk holding the tree root
    int32_t n_child = 0;   // chunks rooted directly below this one
    int32_t height = 0;  // 0 = no child chunks
    int32_t depth = 0;   // 0 = contains the tree root
    int32_t part_row = -1;  // >= 0: a PART chunk (some children of root_node, a split polytomy); its first extra set row
};

// Children per part of a split polytomy (both algorithms), and the set rows a part's partial result takes: Fitch the
// pair (AND, OR); Sankoff 6 counter bits (at most 63 children are counted per part).
constexpr int32_t kPartChildren = 63;  // the most a part may hold (6-bit counters)
constexpr int32_t kPartRowsFitch = 2, kPartRowsSankoff = 6;

struct PlanOptions {
    int32_t chunk_nodes = 64;  // cut a subtree off as its own chunk once it holds at least this many not-yet-cut nodes
    int32_t split_arity = 128; // nodes with MORE children than this are split into parts (0: never)
    int32_t part_children = kPartChildren;  // children per part, 1..kPartChildren (smaller values are for tests)
    int32_t part_rows = kPartRowsFitch;  // rows per partial result: the programs of split nodes depend on the algorithm
};

struct Plan {
    int32_t n_nodes = 0, n_internal = 0, n_leaves = 0, n_rows_used = 0;
    int32_t max_arity = 0;
    int32_t n_parts = 0, part_rows = 0;  // part chunks of split polytomies; rows per partial result (0 without parts)
    int32_t n_set_rows = 0;              // rows of the set arrays: n_internal + n_parts * part_rows
    std::vector<int32_t> int_id;       // [V] internal id or -1
    std::vector<int32_t> int_to_node;  // [I]
    std::vector<int32_t> row_to_node;  // [max row + 1] node id or -1
    std::vector<uint32_t> fwd_prog, bwd_prog;
    std::vector<uint32_t> fwd_stream, bwd_stream;
    std::vector<Chunk> chunks;
    // chunk ids in execution order (dependencies first, longest remaining chain first); lvl_off: chunks per height / depth
    std::vector<int32_t> fwd_order, bwd_order;
    std::vector<int32_t> fwd_lvl_off, bwd_lvl_off;
    int32_t fwd_stack_depth = 0, bwd_stack_depth = 0;
    int64_t edges_leaf = 0, edges_internal = 0;
};

// Validates `tree` and builds the plan.  chunk_nodes: cut a subtree off as its own chunk once it holds at least this
// many not-yet-cut nodes.  The part of the tree left above those chunks is cut again with a threshold 8x smaller, and so
// on: chunks shrink towards the root, where few chunks run side by side and the length of one chunk's walk is the
// critical path.  Returns an empty string or an error description.
std::string build_plan(const pmfs_tree& tree, const PlanOptions& opt, Plan& plan);
inline std::string build_plan(const pmfs_tree& tree, int32_t chunk_nodes, Plan& plan) {
    PlanOptions o;
    o.chunk_nodes = chunk_nodes;
    return build_plan(tree, o, plan);
}

}  // namespace pmfs
