// libpmfs device code, part 1: the two walk kernels and everything they are made of -- work-item scheduling, the producer
// warp with its TMA rings, the per-thread view of an item (DevEnv) that walk.cuh's interpreter runs on, emission.
// Everything is a template on the tile shape CFG (walk_params.hpp: TileShape); one translation unit per shape
// (walk_inst.cuh) instantiates the kernels.
#pragma once

#include <cuda_runtime.h>

#include "../../include/pmfs.h"
#include "plan.hpp"
#include "walk.cuh"
#include "walk_params.hpp"

namespace pmfs {

// 64-word sliding window over a word array that every lane of the warp reads at the same, never decreasing index
// (load streams): lane l holds words base+l and base+32+l in two registers, a read is one shuffle, and the next 32
// words are fetched (one coalesced load per warp) a whole window before they are needed.
struct WarpWords {
    const uint32_t* base;
    int32_t len, wbase;
    uint32_t cur, nxt;
    __device__ __forceinline__ void init(const uint32_t* b, int32_t n, int lane) {
        base = b; len = n; wbase = 0;
        cur = lane < n ? __ldg(b + lane) : 0u;
        nxt = 32 + lane < n ? __ldg(b + 32 + lane) : 0u;
    }
    __device__ __forceinline__ uint32_t get(int32_t i, int lane) {
        while (i >= wbase + 32) {
            cur = nxt;
            wbase += 32;
            const int32_t j = wbase + 32 + lane;
            nxt = j < len ? __ldg(base + j) : 0u;
        }
        return __shfl_sync(0xffffffffu, cur, i - wbase);
    }
};

// Tile shape as a type: NW walking warps + 1 producer warp per CTA, NG column groups per walking thread.
template <int NW_, int NG_>
struct TileCfg {
    static constexpr int NW = NW_, NG = NG_;
    static constexpr int TPB = 32 * NW_;          // walking threads per CTA
    static constexpr int Groups = TPB * NG_;      // column groups per work item (x 32 columns)
    static constexpr int Threads = TPB + 32;      // + the producer warp
    static constexpr int UnitBytes = Groups * 16; // one MSA row segment of the tile
    static constexpr int SetBytes = Groups * 64;  // one node's set for the tile
};
// Operand rings of a CTA: leaf-row units in flight (16 B for each group of the tile) and set slots in flight (64 B for each
// group).  Their depths are powers of two chosen per launch (WalkParams::leaf_ring_log2 / set_ring_log2, pmfs.cu
// ring_depths): bottom-up is mostly leaves, top-down needs one set per internal node; narrow tiles have small units, so
// the same bytes in flight need deeper rings, and on a ladder-like tree -- one chain of dependent nodes per column tile --
// the ring depth IS the number of operands in flight per chain.
constexpr int kMaxLeafRing = 32, kMaxSetRing = 16;
constexpr int kProgWords = 256;           // walk programs up to this many words are staged in shared memory

// ---- mbarrier / bulk-copy (TMA) primitives, sm_90+ PTX ------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {  // never suspends
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// slow path: let the hardware park the thread (suspend-time hint) instead of burning issue slots the producer needs
__device__ __forceinline__ bool mbar_try_park(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity), "r"(20000u)
        : "memory");
    return ok != 0;
}
// A wait that can never be satisfied (a bug in the producer/consumer bookkeeping) must not hang the GPU: after
// WalkParams::spin_limit_ns of waiting (wall clock of the GPU, %globaltimer; default 10 s, pmfs_set_watchdog; 0 = wait
// forever, e.g. under a debugger or on a time-sliced GPU) the kernel traps and the host reports a watchdog error.
__device__ __forceinline__ long long timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return (long long)t;
}
static __device__ __noinline__ void mbar_wait_slow(uint32_t bar, uint32_t parity, long long limit_ns) {
    const long long t0 = timer_ns();
    for (uint32_t it = 1; !mbar_try_park(bar, parity); it++)
        if ((it & 255u) == 0 && limit_ns > 0 && timer_ns() - t0 > limit_ns) __trap();
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, long long limit_ns) {
    if (!mbar_try(bar, parity)) mbar_wait_slow(bar, parity, limit_ns);
}
// -DPMFS_WAITPROF (never in the shipped library): cycles spent per wait kind, summed over warps, printed per launch
#ifdef PMFS_WAITPROF
static __device__ unsigned long long g_waitprof[16];  // (per translation unit = per tile shape)
__device__ __forceinline__ long long wp_now() { return clock64(); }
__device__ __forceinline__ void wp_add(int k, long long t0) {
    if ((threadIdx.x & 31) == 0) atomicAdd(&g_waitprof[k], (unsigned long long)(clock64() - t0));
}
struct WaitProf {
    long long t0;
    int k;
    __device__ __forceinline__ explicit WaitProf(int kind) : t0(wp_now()), k(kind) {}
    __device__ __forceinline__ ~WaitProf() { wp_add(k, t0); }
};
#ifdef PMFS_TRACE_ONLY
#define PMFS_WP(kind)
#else
#define PMFS_WP(kind) WaitProf wp_##kind(kind)
#endif
// per-item timeline: {start ns, end ns, chunk, tile | smid << 16}
constexpr int kTraceCap = 1 << 17;
static __device__ ulonglong4 g_trace[kTraceCap];
static __device__ unsigned int g_trace_n;
__device__ __forceinline__ unsigned long long trace_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void trace_item(unsigned long long t0, int chunk, int tile) {
    if (threadIdx.x != 0) return;
    unsigned int smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    const unsigned int i = atomicAdd(&g_trace_n, 1u);
    if (i < kTraceCap) g_trace[i] = make_ulonglong4(t0, trace_now(), (unsigned long long)chunk, (unsigned long long)tile | ((unsigned long long)smid << 16));
}
#define PMFS_TRACE_BEGIN const unsigned long long trace_t0 = trace_now();
#define PMFS_TRACE_END(chunk, tile) trace_item(trace_t0, chunk, tile);
#else
#define PMFS_TRACE_BEGIN
#define PMFS_TRACE_END(chunk, tile)
#define PMFS_WP(kind)
#endif
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Before the TMA engine may overwrite shared memory that the walking warps have been reading with ordinary loads (a ring
// entry or a program buffer handed back through its `empty` mbarrier): the hand-back orders the walkers' reads before this
// thread's later accesses in the GENERIC proxy only; the refill is a write through the ASYNC proxy, and without this
// fence it can overtake reads that are still in flight.  Seen as one stale operand in ~10^7 (a whole leaf unit of the
// wrong row, almost always in the two back-to-back leaf reads of a cherry), only with one column group per thread,
// where the walkers outrun the producer -- profiles/r2_c_race_hunt.log.
// PMFS_REUSE_MODE (build variants kept for the record of how this was settled, profiles/r2_e_reuse_hunt.log: all three
// give identical records on all of config 4; times are config 2 / config 3's 2 048-column slice):
//   1  fence.proxy.async in the producer before every refill: 0.814 / 26.4 ms -- the unqualified fence also orders the
//      producer's global-memory accesses, which stalls the ring
//   2  the same fence restricted to shared memory (.shared::cta), the state space the hazard is in: 0.730 / 18.7 ms.  DEFAULT
//   3  no fence; the walkers hand an entry back only after its data has ARRIVED in their registers (the hand-back takes a
//      value that depends on the loaded words), so no read of the old contents can be in flight when the refill is
//      issued: 0.741 / 19.0 ms
// Tried and removed: handing an entry back only when the walker asks for its NEXT operand, without a fence -- slower, and
// records still differed (profiles/r2_f_reuse_hunt_mode4.log): consuming the loaded words orders nothing between the proxies.
#ifndef PMFS_REUSE_MODE
#define PMFS_REUSE_MODE 2
#endif
__device__ __forceinline__ void reuse_fence() {
#if PMFS_REUSE_MODE == 1
    asm volatile("fence.proxy.async;" ::: "memory");
#elif PMFS_REUSE_MODE == 2
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#endif
}
// Zero at run time, but only computable once `loaded` -- XOR of one word of every ld.shared of the entry being handed
// back -- has been delivered: added to the address of the hand-back arrive.  (__shfl_sync of a lane's own value.)
__device__ __forceinline__ uint32_t after_loads(uint32_t loaded) {
#if PMFS_REUSE_MODE == 3
    return __shfl_sync(0xffffffffu, loaded, threadIdx.x & 31) ^ loaded;
#else
    return 0u;
#endif
}
// one contiguous global->shared copy by the TMA engine; completion is counted in bytes on `bar`
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}

// Shared-memory layout of one CTA (addresses in the shared window).  Two operand rings, both filled by the producer
// warp with TMA bulk copies and drained by the walking warps: LEAF units (one segment of an MSA row: 16 bytes for each
// of the tile's column groups) and SET slots (one block of the tile-major set array: the 4 quarter-sets of all groups
// of the tile, a single contiguous copy).  Per ring entry: `full` flips when its bytes have landed, `empty` when all
// walking warps have read it.  Entry numbers keep counting across the work items of a persistent CTA, so phases never
// need resetting.
constexpr int kSmemProg = 0;                      // 2 staged programs
constexpr int kSmemDesc = 2 * kProgWords * 4;     // 2 item descriptors (16 B each, 64 B reserved)
constexpr int kSmemLeaf = kSmemDesc + 64;         // leaf ring, then the set ring, then the walk stack (dynamic sizes)
// barriers (static shared array), 8 bytes each
constexpr int kBarLeafFull = 0, kBarLeafEmpty = kMaxLeafRing, kBarSetFull = 2 * kMaxLeafRing, kBarSetEmpty = 2 * kMaxLeafRing + kMaxSetRing,
              kBarProg = 2 * kMaxLeafRing + 2 * kMaxSetRing, kBarDescFull = kBarProg + 2, kBarDescEmpty = kBarProg + 4, kBars = kBarProg + 6;
struct Ring {
    uint32_t smem;       // shared address of the dynamic shared memory
    uint32_t bars;       // shared address of the barrier array
    long long limit_ns;  // watchdog of every wait (WalkParams::spin_limit_ns)
    uint32_t leaf_log2, leaf_mask, set_log2, set_mask;  // ring depths (powers of two)
    uint32_t off_set, off_stack;                        // byte offsets of the set ring and the walk stack
};
// What the producer tells the walkers about a work item (one int4 in shared memory).
struct ItemDesc {
    int32_t chunk;   // -1: no more items
    int32_t tile;
    int32_t plen;    // program length in words
    int32_t parent;  // parent chunk (-1 for the chunk with the tree root)
};

// Lane 0 of the producer warp: the work item AFTER the one whose operands are being streamed.  Its header work (free
// descriptor slot -> program copy -> dependencies met -> descriptor posted) is attempted, without blocking, between
// the ring entries of the current item and only finished with blocking waits once the current stream is done, so the
// walkers normally find the next descriptor and program already there.
struct NextItem {
    int4 a, b;        // packed item records: {chunk, parent, plen, slen}, {prog_off/4, stream_off, need, dep chunk}
    int32_t tile;
    uint32_t ordinal; // how many items this CTA has taken before it (selects descriptor slot and barrier phase)
    bool valid;       // false: no more items, the descriptor only tells the walkers to stop
    bool slot_free, dep_ok, posted, claimed, fetched;
    int32_t item;     // claimed index into the pass's item order
    int32_t t_claim, t_fetch;  // stream entries left when the claim / the record fetch are issued
    __device__ __forceinline__ void begin(uint32_t ord, int32_t slen, int32_t lead) {
        ordinal = ord;
        valid = slot_free = dep_ok = posted = claimed = fetched = false;
        t_claim = min(lead, slen - 1);
        t_fetch = max(t_claim - 2, -1);
    }
    __device__ __forceinline__ void claim(const WalkParams& p) {
        item = atomicAdd(p.next_item, 1);
        claimed = true;
    }
    __device__ __forceinline__ void fetch(const WalkParams& p) {
        if (!claimed) claim(p);
        valid = item < p.n_items;
        tile = 0;
        a = make_int4(-1, -1, 0, 0);
        b = make_int4(0, 0, 0, 0);
        if (valid) {
            const int pos = item / p.ctas_per_chunk;
            tile = item % p.ctas_per_chunk;
            a = __ldg(p.items + 2 * pos);
            b = __ldg(p.items + 2 * pos + 1);
        }
        fetched = true;
    }
};
__device__ __forceinline__ void header_step(const WalkParams& p, const Ring& r, NextItem& h, bool block) {
    const uint32_t slot = h.ordinal & 1u, phase = (h.ordinal >> 1) & 1u;
    if (!h.slot_free) {  // the walkers are done with the item two before this one
        const uint32_t bar = r.bars + (kBarDescEmpty + slot) * 8;
        if (block) {
            PMFS_WP(7);
            mbar_wait(bar, phase ^ 1u, r.limit_ns);
        } else if (!mbar_test(bar, phase ^ 1u)) {
            return;
        }
        h.slot_free = true;
        if (h.valid && h.a.z <= kProgWords) {
            reuse_fence();  // (the walkers read the program of two items ago out of this buffer)
            // (16-byte granularity: programs start at multiples of 4 words and the buffer is padded, see plan.cpp)
            const uint32_t bytes = (uint32_t)((h.a.z + 3) / 4) * 16u;
            mbar_arrive_expect_tx(r.bars + (kBarProg + slot) * 8, bytes);
            bulk_g2s(r.smem + kSmemProg + slot * (kProgWords * 4), p.prog + (int64_t)h.b.x * 4, bytes, r.bars + (kBarProg + slot) * 8);
        }
    }
    if (h.valid && !h.dep_ok) {
        // dependencies: all child chunks (bottom-up) / the parent chunk (top-down), same column tile
        const volatile int* f = p.ready + (int64_t)h.b.w * p.ctas_per_chunk + h.tile;
        if (*f < h.b.z) {
            if (!block) return;
            PMFS_WP(6);
            const long long t0 = timer_ns();
            while (*f < h.b.z) {
                __nanosleep(32);
                if (r.limit_ns > 0 && timer_ns() - t0 > r.limit_ns) __trap();
            }
        }
        __threadfence();
        asm volatile("fence.proxy.async;" ::: "memory");  // data written by other CTAs is read by the TMA engine
        h.dep_ok = true;
    }
    asm volatile("st.shared.v4.s32 [%0], {%1,%2,%3,%4};" ::"r"(r.smem + kSmemDesc + slot * 16), "r"(h.valid ? h.a.x : -1), "r"(h.tile),
                 "r"(h.a.z), "r"(h.a.y)
                 : "memory");
    mbar_arrive(r.bars + (kBarDescFull + slot) * 8);
    h.posted = true;
}

// Producer side of one item: walk the chunk's load stream and keep both rings full.  Called by the whole producer warp
// (the stream words are read warp-wide); lane 0 drives TMA and, in passing, the header of the next item.
template <class CFG>
__device__ __forceinline__ void produce_stream(const WalkParams& p, const Ring& r, WarpWords& w, int32_t slen, int tile,
                                               uint32_t& n_leaf, uint32_t& n_set, NextItem& h) {
    const int lane = threadIdx.x & 31;
    const int64_t g0 = (int64_t)tile * CFG::Groups;
    const uint32_t leaf_bytes = (uint32_t)min((int64_t)CFG::Groups, p.G - g0) * 16u;
    for (int32_t j = 0; j < slen; j++) {
        const uint32_t e = w.get(j, lane);
        if (e & kStreamSet) {
            if (lane == 0) {
                const uint32_t u = n_set & r.set_mask, par = (n_set >> r.set_log2) & 1u;
                PMFS_WP(9);
                mbar_wait(r.bars + (kBarSetEmpty + u) * 8, par ^ 1u, r.limit_ns);  // all walkers are done with the slot's previous set
                reuse_fence();
                // one node's set for this tile: 4 quarters of `tile_w` groups, contiguous (leaf_bytes = tile_w * 16)
                mbar_arrive_expect_tx(r.bars + (kBarSetFull + u) * 8, 4 * leaf_bytes);
                bulk_g2s(r.smem + r.off_set + u * CFG::SetBytes, p.S + (int64_t)(e & ~kStreamSet) * (4 * p.G) + g0 * 4, 4 * leaf_bytes,
                         r.bars + (kBarSetFull + u) * 8);
            }
            n_set++;
        } else {
            if (lane == 0) {
                const uint32_t u = n_leaf & r.leaf_mask, par = (n_leaf >> r.leaf_log2) & 1u;
                PMFS_WP(8);
                mbar_wait(r.bars + (kBarLeafEmpty + u) * 8, par ^ 1u, r.limit_ns);
                reuse_fence();
                mbar_arrive_expect_tx(r.bars + (kBarLeafFull + u) * 8, leaf_bytes);
                bulk_g2s(r.smem + kSmemLeaf + u * CFG::UnitBytes, p.msa + (int64_t)e * p.msa_stride16 + g0, leaf_bytes,
                         r.bars + (kBarLeafFull + u) * 8);
            }
            n_leaf++;
        }
        if (lane == 0) {
            // the next item: claimed `claim_lead` entries before the end of this stream (late, so that an item that
            // is ready is not held by a busy CTA while others idle), its records fetched two entries later, then its
            // header attempted every other entry
            const int32_t left = slen - 1 - j;
            if (left == h.t_claim) h.claim(p);
            else if (left == h.t_fetch) h.fetch(p);
            else if (left < h.t_fetch && !h.posted && (left & 1) == 0) header_step(p, r, h, false);
        }
    }
}

// Per-thread view of one work item for the 4 walking warps: thread t walks column groups tile*256 + t and
// tile*256 + 128 + t.  Operands whose addresses are known from the plan (leaf rows, sets written by earlier work
// items) never go through the walking threads: the producer warp streams them through the shared-memory rings with TMA
// bulk copies, so a walk reads its operands from shared memory and only waits on an mbarrier that has normally
// completed long before.  The walk program itself is staged in shared memory the same way.
// Set arrays are TILE-MAJOR: S[internal id][tile][quarter][groups of the tile] (uint4; 256 groups per tile, fewer in
// the last one), so one node's set for one tile is a single contiguous block of up to 16 KB (one TMA copy) and a
// warp's 128-bit stores are 512-byte segments.  A node's row is exactly 4*G uint4: no padding.
template <class CFG, bool READONLY_SETS>
struct DevEnv {
    static constexpr int kGroups = CFG::NG;
    const WalkParams& p;
    const Ring& ring;
    int64_t g[CFG::NG];      // clamped column groups
    bool active[CFG::NG];
    int64_t gwarp;       // (unclamped) column group of this warp's lane 0, first group slot
    int64_t set_tile;    // uint4 offset of this tile inside one node's set row: tile * 1024
    int32_t tile_w;      // column groups of this tile (256, less for the last tile): stride between the 4 quarters
    int32_t item_tile;   // column tile of this work item
    uint32_t& n_leaf;    // running ring entry numbers of this CTA (same sequences as the producer's)
    uint32_t& n_set;
    uint4* fstk;         // [depth][CFG::NG][4][CFG::TPB]
    uint32_t* bstk;      // [depth][CFG::NG][5][CFG::TPB]
    int sp = 0;
    int lane;
    bool prog_staged;
    const uint32_t* gprog;
    uint32_t sprog;      // shared address of this item's staged program
    uint32_t ltmask;     // bits of a group word whose column is smaller than the column of bit `lane`
    mutable unsigned long long blk_base[CFG::NG];  // lane 0, cell mode: the warp's current output block per group slot
    mutable uint32_t blk_used[CFG::NG], blk_size[CFG::NG];

    __device__ DevEnv(const WalkParams& p_, const Ring& r, const ItemDesc& d, uint32_t slot, uint32_t prog_parity,
                      uint32_t& n_leaf_, uint32_t& n_set_, void* smem_generic)
        : p(p_), ring(r), gwarp((int64_t)d.tile * CFG::Groups + (threadIdx.x & ~31)), set_tile((int64_t)d.tile * (CFG::SetBytes / 16)),
          tile_w((int32_t)min((int64_t)CFG::Groups, p_.G - (int64_t)d.tile * CFG::Groups)),
          item_tile(d.tile),
          n_leaf(n_leaf_), n_set(n_set_), fstk((uint4*)((char*)smem_generic + r.off_stack)),
          bstk((uint32_t*)((char*)smem_generic + r.off_stack)), lane(threadIdx.x & 31), prog_staged(d.plen <= kProgWords),
          gprog(nullptr), sprog(r.smem + kSmemProg + slot * (kProgWords * 4)) {
#pragma unroll
        for (int j = 0; j < CFG::NG; j++) {
            const int64_t gj = (int64_t)d.tile * CFG::Groups + j * CFG::TPB + threadIdx.x;
            active[j] = gj < p_.G;
            g[j] = active[j] ? gj : p_.G - 1;
        }
        // bit q is column 8*(q&3) + (q>>2): smaller residues q&3 come first as a whole, then smaller q>>2 of this lane's own
        ltmask = 0x11111111u * ((1u << (lane & 3)) - 1u) | ((0x11111111u << (lane & 3)) & ((1u << (4 * (lane >> 2))) - 1u));
        PMFS_WP(1);
        if (prog_staged) mbar_wait(r.bars + (kBarProg + slot) * 8, prog_parity, r.limit_ns);
        else gprog = p_.prog + p_.chunk_off[d.chunk];
    }

    __device__ __forceinline__ uint32_t word(const uint32_t*, int32_t pc) const {
        if (prog_staged) {
            uint32_t w;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w) : "r"(sprog + pc * 4));
            return w;
        }
        return __ldg(gprog + pc);  // very long programs (huge polytomies): straight from L1/L2
    }

    // next leaf unit: wait until the TMA copy has landed, read this thread's 16 bytes per group, hand the unit back
    __device__ __forceinline__ void leaf(uint32_t, uint4 (&w)[CFG::NG]) {
        const uint32_t u = n_leaf & ring.leaf_mask, par = (n_leaf >> ring.leaf_log2) & 1u;
        PMFS_WP(2);
        mbar_wait(ring.bars + (kBarLeafFull + u) * 8, par, ring.limit_ns);
        const uint32_t base = ring.smem + kSmemLeaf + u * CFG::UnitBytes + threadIdx.x * 16;
        uint32_t loaded = 0;
#pragma unroll
        for (int j = 0; j < CFG::NG; j++) {
            w[j] = lds128(base + j * (CFG::TPB * 16));
            loaded ^= w[j].x;
        }
        const uint32_t zero = after_loads(loaded);
        __syncwarp();
        if (lane == 0) mbar_arrive(ring.bars + (kBarLeafEmpty + u) * 8 + zero);
        n_leaf++;
    }
    // next set slot (a set written by an earlier work item): 4 quarters x CFG::NG groups
    __device__ __forceinline__ void load_x(uint32_t, Planes16 (&s)[CFG::NG]) {
        const uint32_t u = n_set & ring.set_mask, par = (n_set >> ring.set_log2) & 1u;
        PMFS_WP(3);
        mbar_wait(ring.bars + (kBarSetFull + u) * 8, par, ring.limit_ns);
        const uint32_t base = ring.smem + ring.off_set + u * CFG::SetBytes + threadIdx.x * 16;
        uint32_t loaded = 0;
#pragma unroll
        for (int j = 0; j < CFG::NG; j++)
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const uint4 v = lds128(base + (i * tile_w + j * CFG::TPB) * 16);
                s[j].p[4 * i] = v.x; s[j].p[4 * i + 1] = v.y; s[j].p[4 * i + 2] = v.z; s[j].p[4 * i + 3] = v.w;
                loaded ^= v.x;
            }
        const uint32_t zero = after_loads(loaded);
        __syncwarp();
        if (lane == 0) mbar_arrive(ring.bars + (kBarSetEmpty + u) * 8 + zero);
        n_set++;
    }
    // direct access to the set arrays (sets of this work item itself; Sankoff D1 on the slow path)
    __device__ __forceinline__ int64_t set_index(uint32_t id, int j, int i) const {
        return (int64_t)id * (4 * p.G) + set_tile + i * tile_w + j * CFG::TPB + threadIdx.x;
    }
    __device__ __forceinline__ void load_arr(const uint4* base, uint32_t id, Planes16 (&s)[CFG::NG]) const {
#pragma unroll
        for (int j = 0; j < CFG::NG; j++) {
            if (!active[j]) {  // padding lane of the last tile: its slot lies outside this tile's block
                zero16(s[j]);
                continue;
            }
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const uint4* q = base + set_index(id, j, i);
                const uint4 v = READONLY_SETS ? __ldg(q) : *q;
                s[j].p[4 * i] = v.x; s[j].p[4 * i + 1] = v.y; s[j].p[4 * i + 2] = v.z; s[j].p[4 * i + 3] = v.w;
            }
        }
    }
    __device__ __forceinline__ void store_arr(uint4* base, uint32_t id, const Planes16 (&s)[CFG::NG]) const {
#pragma unroll
        for (int j = 0; j < CFG::NG; j++) {
            if (!active[j]) continue;
#pragma unroll
            for (int i = 0; i < 4; i++)
                base[set_index(id, j, i)] = make_uint4(s[j].p[4 * i], s[j].p[4 * i + 1], s[j].p[4 * i + 2], s[j].p[4 * i + 3]);
        }
    }
    __device__ __forceinline__ void load_s(uint32_t id, Planes16 (&s)[CFG::NG]) const { load_arr(p.S, id, s); }
    __device__ __forceinline__ void load_d(uint32_t id, Planes16 (&s)[CFG::NG]) const { load_arr(p.D, id, s); }
    __device__ __forceinline__ void store_s(uint32_t id, const Planes16 (&s)[CFG::NG]) const { store_arr(p.S, id, s); }
    __device__ __forceinline__ void store_d(uint32_t id, const Planes16 (&s)[CFG::NG]) const { store_arr(p.D, id, s); }

    __device__ __forceinline__ void push(const Planes16 (&s)[CFG::NG]) {
        uint4* q = fstk + (sp * CFG::NG * 4) * CFG::TPB + threadIdx.x;
#pragma unroll
        for (int j = 0; j < CFG::NG; j++)
#pragma unroll
            for (int i = 0; i < 4; i++)
                q[(j * 4 + i) * CFG::TPB] = make_uint4(s[j].p[4 * i], s[j].p[4 * i + 1], s[j].p[4 * i + 2], s[j].p[4 * i + 3]);
        sp++;
    }
    __device__ __forceinline__ void pop(Planes16 (&s)[CFG::NG]) {
        sp--;
        const uint4* q = fstk + (sp * CFG::NG * 4) * CFG::TPB + threadIdx.x;
#pragma unroll
        for (int j = 0; j < CFG::NG; j++)
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const uint4 v = q[(j * 4 + i) * CFG::TPB];
                s[j].p[4 * i] = v.x; s[j].p[4 * i + 1] = v.y; s[j].p[4 * i + 2] = v.z; s[j].p[4 * i + 3] = v.w;
            }
    }
    __device__ __forceinline__ void pushp(const Code4 (&c)[CFG::NG], const uint32_t (&pv)[CFG::NG]) {
        uint32_t* q = bstk + (sp * CFG::NG * 5) * CFG::TPB + threadIdx.x;
#pragma unroll
        for (int j = 0; j < CFG::NG; j++) {
#pragma unroll
            for (int i = 0; i < 4; i++) q[(j * 5 + i) * CFG::TPB] = c[j].b[i];
            q[(j * 5 + 4) * CFG::TPB] = pv[j];
        }
        sp++;
    }
    __device__ __forceinline__ void popp(Code4 (&c)[CFG::NG], uint32_t (&pv)[CFG::NG]) {
        sp--;
        const uint32_t* q = bstk + (sp * CFG::NG * 5) * CFG::TPB + threadIdx.x;
#pragma unroll
        for (int j = 0; j < CFG::NG; j++) {
#pragma unroll
            for (int i = 0; i < 4; i++) c[j].b[i] = q[(j * 5 + i) * CFG::TPB];
            pv[j] = q[(j * 5 + 4) * CFG::TPB];
        }
    }
    // The parent state of a child chunk, and -- top-down dependencies are per chunk, not per finished item -- the go-ahead
    // for it: the child chunk's tile can start while this item is still walking the rest of its own nodes.
    __device__ __forceinline__ void store_x(uint32_t chunk, const Code4 (&c)[CFG::NG], const uint32_t (&pv)[CFG::NG]) const {
#pragma unroll
        for (int j = 0; j < CFG::NG; j++) {
            if (!active[j]) continue;
            p.xcode[(int64_t)chunk * p.G + g[j]] = make_uint4(c[j].b[0], c[j].b[1], c[j].b[2], c[j].b[3]);
            p.xvalid[(int64_t)chunk * p.G + g[j]] = pv[j];
        }
        asm volatile("bar.sync 1, %0;" ::"n"(CFG::TPB) : "memory");  // all four walking warps have stored their part
        if (threadIdx.x == 0) {
            __threadfence();
            atomicAdd(p.ready + (int64_t)chunk * p.ctas_per_chunk + item_tile, 1);
        }
    }
    __device__ __forceinline__ bool any(bool x) const { return __any_sync(0xffffffffu, x); }

    // Stream compaction of one node's mutations.  Called by all 32 lanes (the caller branches on a warp vote); the slots
    // [base, base + total) were reserved by the caller, `base` is advanced past them.  A warp prefix sum over the lanes'
    // record counts gives every lane the place of its own records (lane order = column-group order, so ranks are column
    // order).  Lanes with a few mutations -- the common case away from blocks of N: single substitutions and indels
    // scattered over the lanes -- write their own records, one loop over their set bits in column order.  Lanes with MANY
    // (a 32-column block of N, a long insertion) are served by the whole warp instead: mask and planes are broadcast and
    // each lane writes the record of "its" bit, so that such a block costs ~40 instructions, with coalesced stores.
    // (Before the per-lane path every lane with mutations was served cooperatively: on the bacterial-shaped config 4,
    // where the average call has seven lanes with one mutation each, emission was most of the top-down pass.)
    static constexpr uint32_t kDenseLane = 4;
    __device__ __forceinline__ unsigned long long make_record(uint32_t node, unsigned long long pos, uint32_t cg, uint32_t cp) const {
        // parent gap -> insertion; node gap -> deletion; else substitution (fitchSankoff.cpp:139-167, :683-698)
        const uint32_t type = cp == 0 ? PMFS_NI : (cg == 0 ? PMFS_ND : PMFS_NS);
        const unsigned long long key = ((unsigned long long)node << p.pos_bits) | pos;
        return (key << kPayloadBits) | type | (cg << 2) | (cp << 6);
    }
    __device__ void emit1(uint32_t node, unsigned long long* out, unsigned long long& base, uint32_t m, const Code4& G,
                          const Code4& P, int64_t gw) const {
        const unsigned full = 0xffffffffu;
        const unsigned long long col0 = (unsigned long long)(p.pos_offset + p.col_base + gw * 32);  // first column of lane 0's group
        const unsigned long long pos0 = col0 + (unsigned long long)column_of_bit(lane);
        uint32_t srcs = __ballot_sync(full, m != 0);
        if (__popc(srcs) <= 2) {
            // one or two lanes with mutations (a block of N at a leaf, a lone substitution: nearly every call on
            // SARS-CoV-2-shaped data): serve them cooperatively right away, without the prefix sum
            while (srcs) {
                const int src = __ffs(srcs) - 1;
                srcs &= srcs - 1;
                const uint32_t mm = __shfl_sync(full, m, src);
                uint32_t cg = 0, cp = 0;
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    cg |= ((__shfl_sync(full, G.b[i], src) >> lane) & 1u) << i;
                    cp |= ((__shfl_sync(full, P.b[i], src) >> lane) & 1u) << i;
                }
                if ((mm >> lane) & 1u) {
                    const unsigned long long idx = base + __popc(mm & ltmask);  // rank in COLUMN order
                    if (idx < p.out_cap) out[idx] = make_record(node, pos0 + (unsigned long long)src * 32, cg, cp);
                }
                base += __popc(mm);
            }
            return;
        }
        const uint32_t cnt = __popc(m);
        uint32_t incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(full, incl, d);
            if (lane >= d) incl += t;
        }
        const uint32_t excl = incl - cnt, total = __shfl_sync(full, incl, 31);
        if (cnt != 0 && cnt <= kDenseLane) {
            unsigned long long idx = base + excl;
            const unsigned long long mine = col0 + (unsigned long long)lane * 32;
#pragma unroll 1
            for (int r = 0; r < 4; r++) {  // bit q is column 8*(q&3) + (q>>2): residue by residue, ascending q inside
                uint32_t mr = m & (0x11111111u << r);
                while (mr) {
                    const int q = __ffs(mr) - 1;
                    mr &= mr - 1;
                    const uint32_t cg = ((G.b[0] >> q) & 1u) | (((G.b[1] >> q) & 1u) << 1) | (((G.b[2] >> q) & 1u) << 2) | (((G.b[3] >> q) & 1u) << 3);
                    const uint32_t cp = ((P.b[0] >> q) & 1u) | (((P.b[1] >> q) & 1u) << 1) | (((P.b[2] >> q) & 1u) << 2) | (((P.b[3] >> q) & 1u) << 3);
                    if (idx < p.out_cap) out[idx] = make_record(node, mine + (unsigned long long)(8 * r + (q >> 2)), cg, cp);
                    idx++;
                }
            }
        }
        srcs = __ballot_sync(full, cnt > kDenseLane);
        while (srcs) {
            const int src = __ffs(srcs) - 1;
            srcs &= srcs - 1;
            const uint32_t mm = __shfl_sync(full, m, src);
            const unsigned long long at = base + __shfl_sync(full, excl, src);
            uint32_t cg = 0, cp = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                cg |= ((__shfl_sync(full, G.b[i], src) >> lane) & 1u) << i;
                cp |= ((__shfl_sync(full, P.b[i], src) >> lane) & 1u) << i;
            }
            if ((mm >> lane) & 1u) {
                const unsigned long long idx = at + __popc(mm & ltmask);  // rank in COLUMN order
                if (idx < p.out_cap) out[idx] = make_record(node, pos0 + (unsigned long long)src * 32, cg, cp);
            }
        }
        base += total;
    }
    // Two output layouts.  Regions per 32 column groups (single-GPU runs): the records of one (node, region) pair come
    // from exactly one call, are written contiguously in column order, and regions are laid out in column order -- so a
    // STABLE sort on the node bits alone yields the (node,pos) order.  The regions are further split by the LOW node bits
    // (bucket-major layout): that is the first pass of an LSD radix sort done for free while emitting, and the stable
    // sort only has to cover the remaining high node bits (one radix pass for up to 2^16 nodes instead of four on the
    // whole key).  Regions per owning rank (column shards): see pmfs_set_shard.
    //
    // Cell mode (single-GPU runs, see order_cells_kernel): the ordering kernel is told where every (node, region) cell
    // starts, so a region need not be dense.  Each warp therefore reserves space in BLOCKS of kEmitBlock records -- one
    // atomic per work item and column-group slot, issued when the item starts and not waited for until the first node
    // that has mutations -- and hands out the block to its nodes from a register.  The round trip of a per-node atomic
    // was the longest latency on the top-down path, which is what the chain of chunks below the root is made of.
    static constexpr uint32_t kEmitBlock = 256;
    __device__ __forceinline__ void reserve_blocks() const {
        if (lane != 0) return;
#pragma unroll
        for (int j = 0; j < CFG::NG; j++) {
            blk_used[j] = blk_size[j] = 0;
            if (gwarp + j * CFG::TPB < p.G) {
                blk_base[j] = atomicAdd(p.out_count + p.region_base + ((gwarp + j * CFG::TPB) >> 5), (unsigned long long)kEmitBlock);
                blk_size[j] = kEmitBlock;
            }
        }
    }
    __device__ void emit_cells(uint32_t node, const uint32_t (&m)[CFG::NG], const Code4 (&G)[CFG::NG], const Code4 (&P)[CFG::NG]) const {
#pragma unroll
        for (int j = 0; j < CFG::NG; j++) {
            if (!__any_sync(0xffffffffu, m[j] != 0)) continue;
            const uint32_t total = __reduce_add_sync(0xffffffffu, (uint32_t)__popc(m[j]));
            const int64_t region = p.region_base + ((gwarp + j * CFG::TPB) >> 5);
            unsigned long long base = 0;
            if (lane == 0) {
                if (total > blk_size[j] - blk_used[j]) {  // block used up (or a node with more records than a block holds)
                    blk_size[j] = total > kEmitBlock ? total : kEmitBlock;
                    blk_base[j] = atomicAdd(p.out_count + region, (unsigned long long)blk_size[j]);
                    blk_used[j] = 0;
                }
                base = blk_base[j] + blk_used[j];
                blk_used[j] += total;
                const int64_t cell = (int64_t)node * p.n_regions + region;  // (a cell is emitted at most once per run)
                p.cell_count[cell] = total;
                p.cell_start[cell] = (uint32_t)base;
            }
            base = __shfl_sync(0xffffffffu, base, 0);
            emit1(node, p.out_keys + (unsigned long long)region * p.out_cap, base, m[j], G[j], P[j], gwarp + j * CFG::TPB);
        }
    }
    __device__ void emit(uint32_t node, const uint32_t (&m)[CFG::NG], const Code4 (&G)[CFG::NG], const Code4 (&P)[CFG::NG]) const {
        if (p.cell_count) {
            emit_cells(node, m, G, P);
            return;
        }
        if (p.n_regions > 0) {
            int64_t own_base = 0;  // column shards: regions are owner-major (owner of this node's list, uniform in the warp)
#pragma unroll 1  // (unrolled 16-fold this loop was a fifth of the kernel's code)
            for (int r = 1; r < p.n_owners; r++) own_base += (int32_t)node >= p.owner_bound[r] ? p.n_regions : 0;
#pragma unroll
            for (int j = 0; j < CFG::NG; j++) {
                if (!__any_sync(0xffffffffu, m[j] != 0)) continue;
                const uint32_t total = __reduce_add_sync(0xffffffffu, (uint32_t)__popc(m[j]));
                const int64_t region = own_base + (int64_t)(node & p.bucket_mask) * p.n_regions + p.region_base + ((gwarp + j * CFG::TPB) >> 5);
                unsigned long long base = 0;
                if (lane == 0) base = atomicAdd(p.out_count + region, (unsigned long long)total);
                base = __shfl_sync(0xffffffffu, base, 0);
                emit1(node, p.out_keys + (unsigned long long)region * p.out_cap, base, m[j], G[j], P[j], gwarp + j * CFG::TPB);
            }
            return;
        }
        uint32_t cnt = 0;
#pragma unroll
        for (int j = 0; j < CFG::NG; j++) cnt += __popc(m[j]);
        const uint32_t total = __reduce_add_sync(0xffffffffu, cnt);
        int own = 0;  // owner of this node's list (column shards); uniform across the warp
#pragma unroll 1
        for (int r = 1; r < p.n_owners; r++) own += (int32_t)node >= p.owner_bound[r];
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(p.out_count + own, (unsigned long long)total);
        base = __shfl_sync(0xffffffffu, base, 0);
        unsigned long long* out = p.out_keys + (unsigned long long)own * p.out_cap;
#pragma unroll
        for (int j = 0; j < CFG::NG; j++) emit1(node, out, base, m[j], G[j], P[j], gwarp + j * CFG::TPB);
    }
    __device__ __forceinline__ void emit_internal(uint32_t id, const uint32_t (&m)[CFG::NG], const Code4 (&G)[CFG::NG], const Code4 (&P)[CFG::NG]) const {
        emit((uint32_t)__ldg(p.int_to_node + id), m, G, P);
    }
    __device__ __forceinline__ void emit_leaf(uint32_t row, const uint32_t (&m)[CFG::NG], const Code4 (&G)[CFG::NG], const Code4 (&P)[CFG::NG]) const {
        emit((uint32_t)__ldg(p.row_to_node + row), m, G, P);
    }
};

__device__ __forceinline__ uint32_t real_column_mask(int64_t g, int64_t tile_cols) {
    int64_t n = tile_cols - g * 32;
    if (n >= 32) return ~0u;
    uint32_t m = 0;
#pragma unroll 1  // once per work item and only in the last column group: keep it out of the instruction cache
    for (int q = 0; q < 32; q++)
        if (column_of_bit(q) < n) m |= 1u << q;
    return m;
}

// Work item = (chunk, column tile of 256 groups).  One launch per pass: persistent CTAs claim items from a global
// counter in plan order (bottom-up: chunk height ascending; top-down: chunk depth ascending), so everything an item
// depends on has a smaller index and was claimed earlier by a CTA that is already running -- waiting on it cannot
// deadlock, whatever the number of resident CTAs.  Dependencies are per column tile (tile t of a chunk needs only tile
// t of its child chunks / its parent chunk), tracked by one counter per (chunk, tile); there is no grid-wide barrier
// and no per-level launch, so narrow levels near the root overlap with wide ones instead of leaving the GPU idle.
//
// Inside a CTA the producer warp runs AHEAD of the four walking warps, across item boundaries: while they finish item
// n it has already claimed item n+1, waited for its dependencies, staged its program and started filling the rings
// with its operands, so the walkers go from one item to the next without seeing the claim / dependency / staging
// latencies.  Item descriptors and programs are double buffered; all hand-overs are mbarriers, the kernels contain no
// CTA-wide barrier after set-up.

template <class CFG>
__device__ __forceinline__ Ring ring_setup(void* smem, unsigned long long* bars, const WalkParams& p) {
    Ring r;
    r.smem = (uint32_t)__cvta_generic_to_shared(smem);
    r.bars = (uint32_t)__cvta_generic_to_shared(bars);
    r.limit_ns = p.spin_limit_ns;
    r.leaf_log2 = (uint32_t)p.leaf_ring_log2;
    r.leaf_mask = (1u << r.leaf_log2) - 1u;
    r.set_log2 = (uint32_t)p.set_ring_log2;
    r.set_mask = (1u << r.set_log2) - 1u;
    r.off_set = kSmemLeaf + (CFG::UnitBytes << r.leaf_log2);
    r.off_stack = r.off_set + (CFG::SetBytes << r.set_log2);
    if (threadIdx.x == 0) {
        for (uint32_t i = 0; i <= r.leaf_mask; i++) {
            mbar_init(r.bars + (kBarLeafFull + i) * 8, 1);           // one arrive.expect_tx by the producer
            mbar_init(r.bars + (kBarLeafEmpty + i) * 8, CFG::NW);    // one arrive per walking warp
        }
        for (uint32_t i = 0; i <= r.set_mask; i++) {
            mbar_init(r.bars + (kBarSetFull + i) * 8, 1);
            mbar_init(r.bars + (kBarSetEmpty + i) * 8, CFG::NW);
        }
        for (int i = 0; i < 2; i++) {
            mbar_init(r.bars + (kBarProg + i) * 8, 1);
            mbar_init(r.bars + (kBarDescFull + i) * 8, 1);
            mbar_init(r.bars + (kBarDescEmpty + i) * 8, CFG::NW);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    return r;
}

// Producer warp: claim -> item record -> (program copy, stream window) -> dependency -> descriptor -> operand stream,
// item after item.  Everything that does not need the dependency is started before it is polled, the item record is
// one 32-byte load, and the claim for the NEXT item is issued before this one is produced, so that per item only the
// record load and the dependency poll remain on the producer's critical path.
template <class CFG>
__device__ __forceinline__ void producer_loop(const WalkParams& p, const Ring& r) {
    const int lane = threadIdx.x & 31;
    uint32_t n_leaf = 0, n_set = 0;
    NextItem h = {};
    if (lane == 0) {
        h.begin(0, 1, 0);
        h.fetch(p);
        header_step(p, r, h, true);
    }
    for (uint32_t n = 0;; n++) {
        // item n: descriptor posted; broadcast what the stream needs
        const int32_t chunk = __shfl_sync(0xffffffffu, h.valid ? h.a.x : -1, 0);
        if (chunk < 0) return;
        const int32_t tile = __shfl_sync(0xffffffffu, h.tile, 0);
        const int32_t slen = __shfl_sync(0xffffffffu, h.a.w, 0);
        const int32_t stream_off = __shfl_sync(0xffffffffu, h.b.y, 0);
        WarpWords w;
        w.init(p.stream + stream_off, slen, lane);
        if (lane == 0) h.begin(n + 1, slen, p.claim_lead);
        produce_stream<CFG>(p, r, w, slen, tile, n_leaf, n_set, h);
        if (lane == 0 && !h.posted) {
            if (!h.fetched) h.fetch(p);
            header_step(p, r, h, true);
        }
    }
}

// Walking warps: take the next descriptor; returns false when there is no more work.
__device__ __forceinline__ bool next_item(const Ring& r, uint32_t n, ItemDesc& d) {
    const uint32_t slot = n & 1u, phase = (n >> 1) & 1u;
    {
        PMFS_WP(0);
        mbar_wait(r.bars + (kBarDescFull + slot) * 8, phase, r.limit_ns);
    }
    asm volatile("ld.shared.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(d.chunk), "=r"(d.tile), "=r"(d.plen), "=r"(d.parent)
                 : "r"(r.smem + kSmemDesc + slot * 16)
                 : "memory");
    return d.chunk >= 0;
}
// Item finished: publish it to dependants (all walkers' stores first), then free the descriptor/program slot.
template <class CFG>
__device__ __forceinline__ void finish_item(int* counter, uint32_t desc_empty_bar) {
    if (counter) {  // (uniform; the top-down pass publishes per child chunk in store_x and has nothing to say here)
        {
            PMFS_WP(4);
            asm volatile("bar.sync 1, %0;" ::"n"(CFG::TPB) : "memory");  // the four walking warps only
        }
        if (threadIdx.x == 0) {
            __threadfence();
            atomicAdd(counter, 1);
        }
    }
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive(desc_empty_bar);
}

template <class CFG, int ALGO, bool POLY>
__global__ void __launch_bounds__(CFG::Threads) fwd_kernel(const WalkParams p) {
    extern __shared__ uint4 smem[];
    __shared__ __align__(8) unsigned long long bars[kBars];
    const Ring ring = ring_setup<CFG>(smem, bars, p);
    if (threadIdx.x >= CFG::TPB) {
        PMFS_WP(10);
        producer_loop<CFG>(p, ring);
        return;
    }
    PMFS_WP(5);
    uint32_t n_leaf = 0, n_set = 0, staged[2] = {0, 0};
    ItemDesc d;
    for (uint32_t n = 0; next_item(ring, n, d); n++) {
        const uint32_t slot = n & 1u;
        PMFS_TRACE_BEGIN
        {
            DevEnv<CFG, false> env(p, ring, d, slot, staged[slot] & 1u, n_leaf, n_set, smem);
            Code4 forced[CFG::NG];
            const Code4* fp = nullptr;
            if (ALGO == kAlgoFitch && p.forced_planes) {
#pragma unroll
                for (int j = 0; j < CFG::NG; j++) {
                    const uint4 v = __ldg(p.forced_planes + env.g[j]);
                    forced[j].b[0] = v.x; forced[j].b[1] = v.y; forced[j].b[2] = v.z; forced[j].b[3] = v.w;
                }
                fp = forced;
            }
            forward_walk<ALGO, POLY>(env, nullptr, d.plen, fp);
        }
        if (d.plen <= kProgWords) staged[slot]++;
        finish_item<CFG>(d.parent >= 0 ? p.ready + (int64_t)d.parent * p.ctas_per_chunk + d.tile : nullptr,
                    ring.bars + (kBarDescEmpty + slot) * 8);
        PMFS_TRACE_END(d.chunk, d.tile)
    }
}

// Resident CTAs the top-down kernel is compiled for: 96 registers per thread for Fitch (4 CTAs of the 160-thread shape
// share an SM: +13 % over 3), 128 for Sankoff, whatever the number of threads of the shape.  The 2x1 shape -- what the
// 2 048-column batches of an 8 M-taxon tree on one GPU run -- fits 64 registers without spilling (it takes 86 when left
// alone): top-down pass per batch 10.2 ms at 6 CTAs per SM (18 warps), 8.9 at 8 (80 registers), 8.4 at 10 (64 registers
// and a two-slot set ring, which is what lets them in: pmfs.cu ring_depths).  Not the 1x1 shape: 12 CTAs of two warps
// are slower than 10 (5.6 -> 6.0 ms; every CTA brings a producer warp and its rings).  profiles/r2_v_regs80.log,
// profiles/r2_x_regs64.log.
#ifndef PMFS_R21  // registers per thread the 2x1 Fitch instance is compiled for (build variants: -DPMFS_R21=80, 96)
#define PMFS_R21 64
#endif
template <class CFG, int ALGO>
constexpr int bwd_min_ctas() {
    // (registers are per SM sub-partition: 16 384 each, warps are spread over the four of them)
    constexpr int regs = ALGO == kAlgoFitch ? (CFG::NW == 2 && CFG::NG == 1 ? PMFS_R21 : 96) : 128;
    return 4 * (16384 / (regs * 32)) / (CFG::Threads / 32);
}
template <class CFG, int ALGO>
__global__ void __launch_bounds__(CFG::Threads, bwd_min_ctas<CFG, ALGO>()) bwd_kernel(const WalkParams p) {
    extern __shared__ uint4 smem[];
    __shared__ __align__(8) unsigned long long bars[kBars];
    const Ring ring = ring_setup<CFG>(smem, bars, p);
    if (threadIdx.x >= CFG::TPB) {
        PMFS_WP(10);
        producer_loop<CFG>(p, ring);
        return;
    }
    PMFS_WP(5);
    uint32_t n_leaf = 0, n_set = 0, staged[2] = {0, 0};
    ItemDesc d;
    for (uint32_t n = 0; next_item(ring, n, d); n++) {
        const uint32_t slot = n & 1u;
        uint32_t err;
        PMFS_TRACE_BEGIN
        {
            DevEnv<CFG, true> env(p, ring, d, slot, staged[slot] & 1u, n_leaf, n_set, smem);
            if (p.cell_count) env.reserve_blocks();
            Code4 P[CFG::NG], forced[CFG::NG];
            uint32_t pv[CFG::NG], colmask[CFG::NG];
            const Code4* fp = nullptr;
#pragma unroll
            for (int j = 0; j < CFG::NG; j++) {
                pv[j] = ~0u;
                if (d.chunk == 0) {
                    const uint4 v = __ldg(p.cons_planes + env.g[j]);
                    P[j].b[0] = v.x; P[j].b[1] = v.y; P[j].b[2] = v.z; P[j].b[3] = v.w;
                    if (ALGO == kAlgoSankoff && p.forced_planes) {
                        const uint4 f = __ldg(p.forced_planes + env.g[j]);
                        forced[j].b[0] = f.x; forced[j].b[1] = f.y; forced[j].b[2] = f.z; forced[j].b[3] = f.w;
                        fp = forced;
                    }
                } else {  // written by another SM during this launch: read through L2
                    const uint4 v = __ldcg(p.xcode + (int64_t)d.chunk * p.G + env.g[j]);
                    P[j].b[0] = v.x; P[j].b[1] = v.y; P[j].b[2] = v.z; P[j].b[3] = v.w;
                    pv[j] = __ldcg(p.xvalid + (int64_t)d.chunk * p.G + env.g[j]);
                }
                colmask[j] = env.active[j] ? real_column_mask(env.g[j], p.tile_cols) : 0u;
            }
            err = backward_walk<ALGO>(env, nullptr, d.plen, P, pv, fp, colmask);
        }
        if (err) atomicOr(p.err_flag, 1u);
        if (d.plen <= kProgWords) staged[slot]++;
        finish_item<CFG>(nullptr, ring.bars + (kBarDescEmpty + slot) * 8);  // (child chunks were released by store_x)
        PMFS_TRACE_END(d.chunk, d.tile)
    }
}

}  // namespace pmfs
