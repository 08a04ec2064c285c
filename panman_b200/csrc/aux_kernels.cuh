// libpmfs device code, part 2: everything around the walks -- code planes, ordering of the records (cell matrix and
// radix paths), device-side <=6-run packing, character -> code conversion, the synthetic alignment generator.
// Included by pmfs.cu only, after walk_kernels.cuh.
#pragma once

namespace pmfs {

// codes[col_base .. col_base+tile_cols) -> planes[G] in the bit order of transpose_nibbles
// (also clears up to two small word arrays -- the step's counters -- so that they need no launches of their own)
__global__ void prep_planes_kernel(const uint8_t* __restrict__ codes, int64_t tile_cols, int64_t G, uint4* __restrict__ planes,
                                   uint32_t* zero_a = nullptr, int64_t n_a = 0, uint32_t* zero_b = nullptr, int64_t n_b = 0) {
    int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (int64_t i = g; i < n_a; i += (int64_t)gridDim.x * blockDim.x) zero_a[i] = 0;
    for (int64_t i = g; i < n_b; i += (int64_t)gridDim.x * blockDim.x) zero_b[i] = 0;
    if (g >= G) return;
    uint32_t b[4] = {0, 0, 0, 0};
    for (int q = 0; q < 32; q++) {
        int64_t col = g * 32 + column_of_bit(q);
        uint32_t c = col < tile_cols ? (codes[col] & 15u) : 0u;
#pragma unroll
        for (int i = 0; i < 4; i++) b[i] |= ((c >> i) & 1u) << q;
    }
    planes[g] = make_uint4(b[0], b[1], b[2], b[3]);
}

__global__ void assemble_kernel(const unsigned long long* __restrict__ recs, int64_t n, int pos_bits, pmfs_mut* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long r = recs[i], k = r >> kPayloadBits;
    pmfs_mut m;
    m.node = (int32_t)(k >> pos_bits);
    m.pos = (int32_t)(k & ((1ull << pos_bits) - 1));
    m.type = (uint8_t)(r & 3);
    m.code = (uint8_t)((r >> 2) & 15);
    m.parent_code = (uint8_t)((r >> 6) & 15);
    m.pad = 0;
    out[i] = m;
}

// ---- device-side <=6-run packing (src/panman.cpp:1439-1455 + NucMut range ctor src/panman.hpp:86-128) ---------------------
// Records are sorted by (node,pos).  A RUN is a maximal stretch of one node with consecutive positions and one type;
// the greedy rule of the reference cuts every run into pieces of 6 from its start, so a packed record starts at every
// record whose index inside its run is a multiple of 6.  run_start[i] = index of the first record of i's run
// (max-scan of "i if a run starts here"), starts are counted by a prefix sum, and each start packs up to 6 records.
__global__ void pack_runstart_kernel(const pmfs_mut* __restrict__ m, int64_t n, int32_t* __restrict__ run_start) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    bool brk = i == 0;
    if (!brk) {
        const pmfs_mut a = m[i - 1], b = m[i];
        brk = a.node != b.node || b.pos != a.pos + 1 || a.type != b.type;
    }
    run_start[i] = brk ? (int32_t)i : 0;
}
__global__ void pack_isstart_kernel(const int32_t* __restrict__ run_start, int64_t n, int32_t* __restrict__ is_start) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) is_start[i] = ((int32_t)i - run_start[i]) % 6 == 0;
}
__global__ void pack_emit_kernel(const pmfs_mut* __restrict__ m, const int32_t* __restrict__ run_start,
                                 const int32_t* __restrict__ out_idx, int64_t n, PackedNucMut* __restrict__ out,
                                 int32_t* __restrict__ out_node) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t rs = run_start[i];
    if (((int32_t)i - rs) % 6 != 0) return;
    int len = 1;
    while (len < 6 && i + len < n && run_start[i + len] == rs) len++;
    PackedNucMut r;
    r.nucPosition = m[i].pos;
    r.nucGapPosition = -1;
    r.primaryBlockId = 0;
    r.secondaryBlockId = -1;
    r.mutInfo = (uint32_t)(uint8_t)((len << 4) + m[i].type);
    r.nucs = 0;
    for (int k = 0; k < len; k++) r.nucs += (uint32_t)m[i + k].code << (4 * (5 - k));
    out[out_idx[i]] = r;
    out_node[out_idx[i]] = m[i].node;
}
// node_off[v] = first packed record of node >= v; node_off[n_nodes] = n_packed
__global__ void pack_offsets_kernel(const int32_t* __restrict__ out_node, int64_t n_packed, int32_t n_nodes, int64_t* __restrict__ node_off) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k > n_packed) return;
    const int32_t lo = k == 0 ? 0 : out_node[k - 1] + 1;
    const int32_t hi = k == n_packed ? n_nodes : out_node[k];
    for (int32_t v = lo; v <= hi; v++) node_off[v] = k;
}

// ---- FASTA characters -> packed 4-bit codes on the device -------------------------------------------------
// getCodeFromNucleotide (src/panman.cpp:83-118): case-sensitive, everything that is not one of the 15 upper-case IUPAC
// letters (including '-', lower case and '\0') is code 0.
__device__ __forceinline__ uint32_t code_of_char(uint32_t ch) {
    switch (ch) {
        case 'A': return 1;  case 'C': return 2;  case 'M': return 3;  case 'G': return 4;
        case 'R': return 5;  case 'S': return 6;  case 'V': return 7;  case 'T': return 8;
        case 'W': return 9;  case 'Y': return 10; case 'H': return 11; case 'K': return 12;
        case 'D': return 13; case 'B': return 14; case 'N': return 15;
        default: return 0;
    }
}
// one thread = 8 columns of one row: 8 characters in, one 32-bit word of nibbles out (column 2j in the low nibble of byte j)
template <bool RAW>  // RAW: the bytes are state codes already (low 4 bits), not alignment characters
__global__ void pack_chars_kernel(const uint8_t* __restrict__ chars, int64_t in_stride, int32_t n_rows, int64_t n_cols,
                                  uint8_t* __restrict__ out, int64_t out_stride) {
    const int64_t words = out_stride / 4;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= words * n_rows) return;
    const int64_t row = i / words, w = i % words;
    const uint8_t* src = chars + row * in_stride + w * 8;
    uint32_t v = 0;
    for (int k = 0; k < 8; k++)
        if (w * 8 + k < n_cols) v |= (RAW ? (uint32_t)src[k] & 15u : code_of_char(src[k])) << (4 * k);
    *(uint32_t*)(out + row * out_stride + w * 4) = v;
}

// sum and maximum of the output-area counters (overflow check and record count without copying the counters to the host)
__global__ void area_stats_kernel(const unsigned long long* __restrict__ count, int64_t n_areas, unsigned long long* __restrict__ stats) {
    unsigned long long sum = 0, mx = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_areas; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long v = count[i];
        sum += v;
        mx = mx > v ? mx : v;
    }
    for (int d = 16; d > 0; d >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, d);
        const unsigned long long o = __shfl_xor_sync(0xffffffffu, mx, d);
        mx = mx > o ? mx : o;
    }
    if ((threadIdx.x & 31) == 0) {
        if (sum) atomicAdd(stats, sum);
        if (mx) atomicMax(stats + 1, mx);
    }
}
// one warp per output area (grid-stride): count[a] records at src + a*cap -> dst + offset[a]
__global__ void compact_areas_kernel(const unsigned long long* __restrict__ src, unsigned long long cap,
                                     const unsigned long long* __restrict__ count, const unsigned long long* __restrict__ offset,
                                     int64_t n_areas, const unsigned long long* __restrict__ max_count,
                                     unsigned long long* __restrict__ dst) {
    if (*max_count > cap) return;  // some area overflowed: the host grows the buffer and runs the passes again
    const int lane = threadIdx.x & 31;
    const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t a = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; a < n_areas; a += n_warps) {
        const unsigned long long n = count[a];
        const unsigned long long* s = src + (unsigned long long)a * cap;
        unsigned long long* d = dst + offset[a];
        unsigned long long i = lane;
        for (; i + 96 < n; i += 128) {  // four loads in flight per lane
            const unsigned long long v0 = s[i], v1 = s[i + 32], v2 = s[i + 64], v3 = s[i + 96];
            d[i] = v0; d[i + 32] = v1; d[i + 64] = v2; d[i + 96] = v3;
        }
        for (; i < n; i += 32) d[i] = s[i];
    }
}

// Ordering without a sort (single-GPU runs).  The top-down pass leaves, for every (node, column region) CELL, how many
// records it emitted and where they start inside the region's buffer; inside a cell the records are contiguous and in
// column order.  The (node,pos) order is therefore cell order, node-major: an exclusive prefix sum over the cell counts
// gives every cell its place in the output and this kernel copies the cells there, turning the packed 8-byte records
// into pmfs_mut on the way.  One warp per 32 consecutive cells: their output range is contiguous, so the warp walks it
// 32 records at a time and each lane finds the cell its record belongs to from a mask of the cell starts in the window.
// stats = {records, largest region count}: on overflow of a region or of `out` nothing is written (the host redoes it).
__global__ void cells_total_kernel(const uint32_t* __restrict__ cnt, const uint32_t* __restrict__ prefix, int64_t cells,
                                   unsigned long long* __restrict__ stats) {
    stats[3] = stats[0];  // record slots reserved (area_stats_kernel's sum: an upper bound that cannot wrap)
    stats[0] = (unsigned long long)prefix[cells - 1] + cnt[cells - 1];
}
__device__ __forceinline__ int kth_set_bit(uint32_t mask, uint32_t k) {  // position of the k-th (from 0) set bit
    int pos = 0;
#pragma unroll
    for (int step = 16; step > 0; step >>= 1) {
        const uint32_t c = __popc((mask >> pos) & ((1u << step) - 1u));
        if (k >= c) {
            pos += step;
            k -= c;
        }
    }
    return pos;
}
// KEYS: the ordered records stay 64-bit words (column shards whose lists are gathered by node range: 8 bytes per record
// cross NVLink and the owner converts, comm.cu); else they become pmfs_mut.
template <bool KEYS>
__global__ void order_cells_kernel(const unsigned long long* __restrict__ keys, unsigned long long cap, uint32_t* __restrict__ cnt,
                                   const uint32_t* __restrict__ start, const uint32_t* __restrict__ prefix, int64_t cells, int32_t W,
                                   int pos_bits, const unsigned long long* __restrict__ stats, unsigned long long out_cap,
                                   void* __restrict__ out) {
    if (stats[1] > cap || stats[0] > out_cap || (stats[3] >> 32)) return;  // (32-bit prefix sums: see cells_total_kernel)
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int64_t stride = (((int64_t)gridDim.x * blockDim.x) >> 5) * 32;
    const unsigned long long pos_mask = (1ull << pos_bits) - 1;
    uint32_t* const out32 = (uint32_t*)out;
    unsigned long long* const out64 = (unsigned long long*)out;
    int64_t c0 = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 32;
    // the three per-cell words of the NEXT block are fetched while the current one is copied
    uint32_t n_nx = 0, s_nx = 0, e_nx = 0;
    if (c0 + lane < cells) {
        n_nx = cnt[c0 + lane];
        s_nx = start[c0 + lane];
        e_nx = prefix[c0 + lane];
    }
    for (; c0 < cells; c0 += stride) {
        const uint32_t n_l = n_nx, s_l = s_nx, e_l = e_nx;
        if (n_l) cnt[c0 + lane] = 0;  // leaves the matrix clear for the next run (no 12-bytes-per-cell memset per step)
        n_nx = 0;
        if (c0 + stride + lane < cells) {
            n_nx = cnt[c0 + stride + lane];
            s_nx = start[c0 + stride + lane];
            e_nx = prefix[c0 + stride + lane];
        }
        const uint32_t nonempty = __ballot_sync(full, n_l != 0);
        if (nonempty == 0) continue;
        const int32_t region = (int32_t)((uint32_t)(c0 + lane) % (uint32_t)W);  // (cells < 2^31)
        const unsigned long long* src = keys + (unsigned long long)region * cap + s_l;
        const uint32_t longest = __reduce_max_sync(full, n_l);
        if (longest <= 24) {
            // the usual case, a handful of records per cell: every lane copies its own cell
            uint32_t* o = out32 + (size_t)e_l * 3;
            unsigned long long r = n_l ? src[0] : 0ull;
            for (uint32_t i = 0; i < n_l; i++, o += 3) {
                const unsigned long long cur = r, key = r >> kPayloadBits;
                const uint32_t pay = (uint32_t)r;
                if (i + 1 < n_l) r = src[i + 1];  // next record in flight while this one is stored
                if (KEYS) {
                    out64[(size_t)e_l + i] = cur;
                } else {
                    o[0] = (uint32_t)(key >> pos_bits);
                    o[1] = (uint32_t)(key & pos_mask);
                    o[2] = (pay & 3) | (((pay >> 2) & 15) << 8) | (((pay >> 6) & 15) << 16);
                }
            }
            continue;
        }
        // some cell is long (a block of N's, an insertion): the warp walks the block's output range 32 records at a time
        const uint32_t e0 = __shfl_sync(full, e_l, 0);
        const uint32_t rel_l = e_l - e0;  // (lanes past the last cell are empty and never looked at)
        const uint32_t total = __reduce_add_sync(full, n_l);
        for (uint32_t t0 = 0; t0 < total; t0 += 32) {
            const bool here = n_l != 0 && rel_l >= t0 && rel_l < t0 + 32;
            const uint32_t marks = __reduce_or_sync(full, here ? 1u << (rel_l - t0) : 0u);
            const uint32_t before = __popc(__ballot_sync(full, n_l != 0 && rel_l < t0));
            const uint32_t t = t0 + lane;
            // the record at output slot t belongs to the k-th non-empty cell of the block (k counted from 0)
            const uint32_t k = before + __popc(marks & ((2u << lane) - 1u)) - 1u;
            const int src_lane = t < total ? kth_set_bit(nonempty, k) : 0;
            const unsigned long long* src_t = (const unsigned long long*)__shfl_sync(full, (unsigned long long)src, src_lane);
            const uint32_t rel_src = __shfl_sync(full, rel_l, src_lane);
            if (t < total) {
                const unsigned long long r = src_t[t - rel_src], key = r >> kPayloadBits;
                if (KEYS) {
                    out64[(size_t)e0 + t] = r;
                } else {
                    uint32_t* o = out32 + ((size_t)e0 + t) * 3;
                    o[0] = (uint32_t)(key >> pos_bits);
                    o[1] = (uint32_t)(key & pos_mask);
                    o[2] = (uint32_t)(r & 3) | ((uint32_t)((r >> 2) & 15) << 8) | ((uint32_t)((r >> 6) & 15) << 16);
                }
            }
        }
    }
}

// Per-node offsets into the ordered records, from the cell prefix sums: noff[v] = first record of node v, noff[V] = n;
// cut[r] = first record of the node range owned by rank r (cut[n_parts] = n).
__global__ void node_offsets_kernel(const uint32_t* __restrict__ prefix, int32_t W, int32_t n_nodes, const unsigned long long* __restrict__ stats,
                                    uint32_t* __restrict__ noff, PartBounds parts, unsigned long long* __restrict__ cut) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n_nodes) noff[v] = prefix[v * W];
    if (v == n_nodes) noff[v] = (uint32_t)stats[0];
    if (v <= parts.n_parts) {
        const int32_t b = parts.bound[v];
        cut[v] = b >= n_nodes ? stats[0] : (unsigned long long)prefix[(int64_t)b * W];
    }
}
// The same from the ordered 64-bit records themselves (radix path: no cell matrix): lower bound of each node id.
__global__ void node_lower_bound_kernel(const unsigned long long* __restrict__ recs, int64_t n, int shift, int32_t n_nodes,
                                        uint32_t* __restrict__ noff, PartBounds parts, unsigned long long* __restrict__ cut) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v > n_nodes) return;
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if ((int64_t)(recs[mid] >> shift) < v) lo = mid + 1;
        else hi = mid;
    }
    noff[v] = (uint32_t)lo;
    for (int r = 0; r <= parts.n_parts; r++)
        if (parts.bound[r] == v || (v == n_nodes && parts.bound[r] > n_nodes)) cut[r] = (unsigned long long)lo;
}

// region r's records (count[r] of them at src + r*cap) -> dst + offset[r]: makes the per-region output contiguous
__global__ void compact_regions_kernel(const unsigned long long* __restrict__ src, unsigned long long cap,
                                       const unsigned long long* __restrict__ count, const unsigned long long* __restrict__ offset,
                                       unsigned long long* __restrict__ dst) {
    const unsigned long long n = count[blockIdx.y], o = offset[blockIdx.y];
    const unsigned long long* s = src + (unsigned long long)blockIdx.y * cap;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x)
        dst[o + i] = s[i];
}

// ------------------------------------------------------------------------------------------------ synthetic alignments
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {  // splitmix64 finaliser
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
    x ^= x >> 27; x *= 0x94d049bb133111ebull;
    x ^= x >> 31;
    return x;
}
__host__ __device__ __forceinline__ uint32_t hash3(uint64_t seed, uint64_t a, uint64_t b) {
    return (uint32_t)(mix64(seed ^ mix64(a * 0x9e3779b97f4a7c15ull + b)) >> 32);
}

// The alignment is simulated level by level down the tree (one launch per tree depth): a node's state is its parent's
// state after the branch events hash3(seed, node, column) decides, so every cell is a pure function of
// (seed, node, column) and any column range can be generated on its own.  Only the previous level's internal-node
// states are kept (packed nibbles, ping-pong buffers).
struct SynthParams {
    uint64_t seed;
    uint32_t t_sub, t_del, t_ins, t_nblock, t_amb, t_acc;  // thresholds on a 32-bit hash
    const int32_t* node;         // [n_level] node ids of this depth
    const int32_t* parent_slot;  // [n_level] slot of the parent in `prev`
    const int32_t* my_slot;      // [n_level] slot in `cur` (internal node) or -(row+1) (leaf with an MSA row), 0x7fffffff: absent leaf
    int32_t n_level;
    int32_t is_root;
    const uint32_t* prev;
    uint32_t* cur;
    int64_t words;               // 8-column words per row
    int32_t n_rows;
    int64_t col_begin, n_cols;
    uint8_t* msa;
    int64_t row_stride;
};

__device__ __forceinline__ uint32_t synth_root(const SynthParams& s, int64_t col, uint32_t& r) {
    r = hash3(s.seed, 0x1000000000ull, (uint64_t)col);
    // root: A .30 C .18 G .20 T .32, or gap in accessory columns
    const uint32_t u = r >> 8;
    if (hash3(s.seed, 0x2000000000ull, (uint64_t)col) < s.t_acc) return 0;
    return u < 5033165u ? 1 : (u < 8053064u ? 2 : (u < 11408507u ? 4 : 8));
}
__device__ __forceinline__ uint32_t synth_branch(const SynthParams& s, uint32_t st, int32_t node, int64_t col) {
    const uint32_t base4[4] = {1, 2, 4, 8};
    const uint32_t h = hash3(s.seed, (uint64_t)node, (uint64_t)col);
    if (st != 0) {
        if (h < s.t_sub) {
            const uint32_t nb = base4[(h >> 3) & 3];
            st = nb == st ? base4[((h >> 3) + 1) & 3] : nb;
        } else if (h - s.t_sub < s.t_del) {
            st = 0;
        }
    } else if (h < s.t_ins) {
        st = base4[(h >> 5) & 3];
    }
    return st;
}
__device__ __forceinline__ uint32_t synth_leaf(const SynthParams& s, uint32_t st, int32_t row, int64_t col) {
    const uint32_t base4[4] = {1, 2, 4, 8};
    if (hash3(s.seed, 0x3000000000ull + (uint64_t)row, (uint64_t)(col >> 5)) < s.t_nblock) st = 15;
    else {
        const uint32_t h = hash3(s.seed, 0x4000000000ull + (uint64_t)row, (uint64_t)col);
        if (h < s.t_amb) {
            const uint32_t amb[10] = {3, 5, 6, 7, 9, 10, 11, 12, 13, 14};
            st = amb[(h >> 7) % 10];
        }
    }
    // guarantee: one designated row per column is never a gap
    if (st == 0 && (uint32_t)(hash3(s.seed, 0x5000000000ull, (uint64_t)col) % (uint32_t)s.n_rows) == (uint32_t)row)
        st = base4[hash3(s.seed, 0x1000000000ull, (uint64_t)col) & 3];
    return st;
}

// one thread = one (node of this depth, 8-column word)
__global__ void synth_level_kernel(const SynthParams s) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s.words * s.n_level) return;
    const int32_t k = (int32_t)(i / s.words);
    const int64_t w = i % s.words;
    const int32_t node = s.node[k], slot = s.my_slot[k];
    if (slot == 0x7fffffff) return;
    const uint32_t parent = s.is_root ? 0u : s.prev[(int64_t)s.parent_slot[k] * s.words + w];
    uint32_t out = 0;
    for (int j = 0; j < 8; j++) {
        const int64_t c = w * 8 + j;
        if (c >= s.n_cols) break;
        const int64_t col = s.col_begin + c;
        uint32_t st, r;
        if (s.is_root) st = synth_root(s, col, r);
        else st = synth_branch(s, (parent >> (4 * j)) & 15u, node, col);
        if (slot < 0) st = synth_leaf(s, st, -slot - 1, col);
        out |= st << (4 * j);
    }
    if (slot >= 0) s.cur[(int64_t)slot * s.words + w] = out;
    else *(uint32_t*)(s.msa + (int64_t)(-slot - 1) * s.row_stride + w * 4) = out;
}

// first non-gap code in row order (src/panman.cpp:1326-1340 with rows in id order)
__global__ void synth_consensus_kernel(const uint8_t* __restrict__ msa, int64_t row_stride, int32_t n_rows, int64_t n_cols,
                                       uint8_t* __restrict__ cons) {
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_cols) return;
    uint32_t code = 0;
    for (int32_t r = 0; r < n_rows && code == 0; r++) code = (msa[(int64_t)r * row_stride + (c >> 1)] >> ((c & 1) * 4)) & 15u;
    cons[c] = (uint8_t)code;
}

}  // namespace pmfs
