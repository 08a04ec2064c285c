// Bit-sliced set algebra and the two walk-program interpreters (bottom-up, top-down).
//
// One thread owns one COLUMN GROUP = 32 alignment columns (one 128-bit load of packed nibbles).  All per-column state
// is bit-sliced: a 16-state set for 32 columns is 16 words ("planes", plane s bit q = state s is in the set of the
// column at bit q); a 4-bit code for 32 columns is 4 words.  Every set operation of the reference then costs about one
// LOP3 per 32 columns per state, with no per-column branches, and the tree walk is the same instruction stream for
// every thread of the grid.
//
// Semantics restated here (paths relative to /root/reference):
//   Fitch   join        src/fitchSankoff.cpp:41-55    and ? and : or  over ALL children (an empty child set kills the AND)
//           resolve     src/fitchSankoff.cpp:96-129   parent & set ? parent : lowest set bit; root: lowest set bit
//           classify    src/fitchSankoff.cpp:131-171
//   Sankoff join        src/fitchSankoff.cpp:359-405  in its exact two-mask form (SURVEY §8 notes): with unit costs
//                       cost[i] = sum_j min_j + #{valid children j : i not in S_j}; S = argmin set, D1 = {cost = min+1}
//           resolve     src/fitchSankoff.cpp:487-531  p in S -> p; p in D1 -> min(p, lowest(S)); else lowest(S)
//           classify    src/fitchSankoff.cpp:676-703
// The functions are __host__ __device__ so that tests/host_emul.cpp can run the very same interpreter on the CPU
// against the oracle (test infrastructure only -- the library has no CPU path).
#pragma once

#include <cstdint>

#include <vector_types.h>

#include "plan.hpp"

#if defined(__CUDACC__)
#define PMFS_HD __host__ __device__ __forceinline__
#else
#define PMFS_HD inline
#endif

namespace pmfs {

struct Planes16 {
    uint32_t p[16];
};
struct Code4 {
    uint32_t b[4];
};

// Bit q of a group's words <-> column 32*g + 8*(q & 3) + (q >> 2)   (see transpose_nibbles)
PMFS_HD int32_t column_of_bit(int32_t q) { return 8 * (q & 3) + (q >> 2); }

// 32 packed nibbles (column 8*i+k of the group in bits [4k,4k+4) of word i) -> 4 code-bit planes.
// Two delta-swap rounds transpose the 4x4 bit matrix (word index x bit-in-nibble) of every nibble position at once,
// so plane j ends up holding code bit j of column 8*i+k at bit 4*k+i.
PMFS_HD void transpose_nibbles(const uint4& w, Code4& c) {
    uint32_t w0 = w.x, w1 = w.y, w2 = w.z, w3 = w.w, t;
    t = ((w0 >> 1) ^ w1) & 0x55555555u; w1 ^= t; w0 ^= t << 1;
    t = ((w2 >> 1) ^ w3) & 0x55555555u; w3 ^= t; w2 ^= t << 1;
    t = ((w0 >> 2) ^ w2) & 0x33333333u; w2 ^= t; w0 ^= t << 2;
    t = ((w1 >> 2) ^ w3) & 0x33333333u; w3 ^= t; w1 ^= t << 2;
    c.b[0] = w0; c.b[1] = w1; c.b[2] = w2; c.b[3] = w3;
}

// code planes -> one-hot set planes (1 << code per column)
PMFS_HD void onehot(const Code4& c, Planes16& s) {
    uint32_t lo[4], hi[4];
    lo[0] = ~c.b[1] & ~c.b[0]; lo[1] = ~c.b[1] & c.b[0]; lo[2] = c.b[1] & ~c.b[0]; lo[3] = c.b[1] & c.b[0];
    hi[0] = ~c.b[3] & ~c.b[2]; hi[1] = ~c.b[3] & c.b[2]; hi[2] = c.b[3] & ~c.b[2]; hi[3] = c.b[3] & c.b[2];
#pragma unroll
    for (int s_ = 0; s_ < 16; s_++) s.p[s_] = hi[s_ >> 2] & lo[s_ & 3];
}

PMFS_HD uint32_t or16(const Planes16& s) {
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) r |= s.p[i];
    return r;
}

PMFS_HD void zero16(Planes16& s) {
#pragma unroll
    for (int i = 0; i < 16; i++) s.p[i] = 0;
}

// s[code] per column: a 16:1 multiplexer tree driven by the code planes
PMFS_HD uint32_t mux16(const Planes16& s, const Code4& c) {
    uint32_t m8[8], m4[4], m2[2];
#pragma unroll
    for (int i = 0; i < 8; i++) m8[i] = (s.p[2 * i + 1] & c.b[0]) | (s.p[2 * i] & ~c.b[0]);
#pragma unroll
    for (int i = 0; i < 4; i++) m4[i] = (m8[2 * i + 1] & c.b[1]) | (m8[2 * i] & ~c.b[1]);
#pragma unroll
    for (int i = 0; i < 2; i++) m2[i] = (m4[2 * i + 1] & c.b[2]) | (m4[2 * i] & ~c.b[2]);
    return (m2[1] & c.b[3]) | (m2[0] & ~c.b[3]);
}

// code of the lowest member of each column's set (0 where the set is empty): a bit-sliced priority encoder
PMFS_HD void lowest(const Planes16& s, uint32_t nz, Code4& l) {
    uint32_t lo8 = s.p[0] | s.p[1] | s.p[2] | s.p[3] | s.p[4] | s.p[5] | s.p[6] | s.p[7];
    uint32_t h[8];
#pragma unroll
    for (int i = 0; i < 8; i++) h[i] = (s.p[i] & lo8) | (s.p[i + 8] & ~lo8);
    uint32_t lo4 = h[0] | h[1] | h[2] | h[3];
    uint32_t q[4];
#pragma unroll
    for (int i = 0; i < 4; i++) q[i] = (h[i] & lo4) | (h[i + 4] & ~lo4);
    uint32_t lo2 = q[0] | q[1];
    uint32_t r0 = (q[0] & lo2) | (q[2] & ~lo2);
    l.b[3] = ~lo8 & nz;
    l.b[2] = ~lo4 & nz;
    l.b[1] = ~lo2 & nz;
    l.b[0] = ~r0 & nz;
}

PMFS_HD uint32_t code_ne(const Code4& a, const Code4& b) {
    return (a.b[0] ^ b.b[0]) | (a.b[1] ^ b.b[1]) | (a.b[2] ^ b.b[2]) | (a.b[3] ^ b.b[3]);
}

// a < b per column (4-bit unsigned)
PMFS_HD uint32_t code_lt(const Code4& a, const Code4& b) {
    uint32_t lt = 0, eq = ~0u;
#pragma unroll
    for (int i = 3; i >= 0; i--) {
        lt |= eq & ~a.b[i] & b.b[i];
        eq &= ~(a.b[i] ^ b.b[i]);
    }
    return lt;
}

PMFS_HD void code_select(uint32_t m, const Code4& a, const Code4& b, Code4& out) {  // m ? a : b
#pragma unroll
    for (int i = 0; i < 4; i++) out.b[i] = (a.b[i] & m) | (b.b[i] & ~m);
}

// Fitch: and ? and : or (src/fitchSankoff.cpp:48-55).  Also the S part of the binary Sankoff join.
PMFS_HD void join_sets(const Planes16& a, const Planes16& b, Planes16& out, uint32_t& any_and) {
    uint32_t any = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) any |= a.p[i] & b.p[i];
#pragma unroll
    for (int i = 0; i < 16; i++) out.p[i] = (a.p[i] & b.p[i]) | ((a.p[i] | b.p[i]) & ~any);
    any_and = any;
}

// Sankoff, two children: D1 = states one step above the minimum (see header comment).
PMFS_HD void join_d1(const Planes16& a, const Planes16& b, uint32_t any_and, Planes16& d1) {
#pragma unroll
    for (int i = 0; i < 16; i++) d1.p[i] = ((a.p[i] ^ b.p[i]) & any_and) | (~(a.p[i] | b.p[i]) & ~any_and);
}

constexpr int kAlgoFitch = 0;
constexpr int kAlgoSankoff = 1;
constexpr int kPolyBits = 28;  // counter width of the any-arity Sankoff join: arity < 2^28 (node ids are < 2^28)

// ---------------------------------------------------------------------------------------------------------------------
// Env concept (device: kernels.cu; host emulation: tests/host_emul.cpp)
//   uint32_t word(const uint32_t* prog, int32_t pc)          program word, same for every thread
//   uint4    leaf(uint32_t row)                               this thread's 32 nibbles of MSA row `row` (next ring entry)
//   void     load_x(uint32_t id, Planes16&)                   set S[id], prefetched (next 4 ring entries)
//   void     load_s(uint32_t id, Planes16&)  / store_s        set array S (Fitch: the only one), direct access
//   void     load_d(uint32_t id, Planes16&)  / store_d        Sankoff D1 array
//   void     push(const Planes16&) / pop(Planes16&)           forward stack
//   void     pushp(const Code4&, uint32_t) / popp(...)        backward stack
//   void     store_x(uint32_t chunk, const Code4&, uint32_t)  parent state for a child chunk
//   bool     any(bool)                                        warp-wide OR (host: identity)
//   void     emit_internal(uint32_t id, uint32_t m, const Code4& g, const Code4& p)
//   void     emit_leaf(uint32_t row, uint32_t m, const Code4& g, const Code4& p)
// ---------------------------------------------------------------------------------------------------------------------

template <class Env>
PMFS_HD void load_operand(Env& env, uint32_t o, Planes16& s) {
    if (o & kOperandSet) {
        if (o & kOperandExt) env.load_x(o & kOperandIdMask, s);  // written by an earlier launch: comes through the ring
        else env.load_s(o & kOperandIdMask, s);                  // written by this thread earlier in this launch
    } else if (o == kAbsentLeaf) {
        zero16(s);
    } else {
        Code4 c;
        transpose_nibbles(env.leaf(o), c);
        onehot(c, s);
    }
}

template <int ALGO, class Env>
PMFS_HD void finish_node(Env& env, uint32_t id, Planes16& S, const Code4* forced) {
    if (ALGO == kAlgoFitch && id == 0 && forced) onehot(*forced, S);  // refState overwrites the root (:45-47)
    env.store_s(id, S);
}

// Any-arity join (unary nodes and polytomies); children are re-read from memory, so nothing is live across them.
template <int ALGO, class Env>
PMFS_HD void poly_join(Env& env, const uint32_t* prog, int32_t pc, uint32_t k, Planes16& S, Planes16& D1) {
    if (ALGO == kAlgoFitch) {
        Planes16 A, O;
#pragma unroll
        for (int i = 0; i < 16; i++) { A.p[i] = ~0u; O.p[i] = 0; }
        for (uint32_t j = 0; j < k; j++) {
            Planes16 B;
            load_operand(env, env.word(prog, pc + (int32_t)j), B);
#pragma unroll
            for (int i = 0; i < 16; i++) { A.p[i] &= B.p[i]; O.p[i] |= B.p[i]; }
        }
        uint32_t any = or16(A);
#pragma unroll
        for (int i = 0; i < 16; i++) S.p[i] = A.p[i] | (O.p[i] & ~any);
    } else {
        // h[s] = number of valid children whose S contains s, as bit-sliced counters; S = argmax h, D1 = {h = max-1}
        int nb = 1;
        while ((k >> nb) != 0) nb++;
        uint32_t h[16][kPolyBits];
        for (int s = 0; s < 16; s++)
            for (int b = 0; b < nb; b++) h[s][b] = 0;
        uint32_t any_valid = 0;
        for (uint32_t j = 0; j < k; j++) {
            Planes16 B;
            load_operand(env, env.word(prog, pc + (int32_t)j), B);
            any_valid |= or16(B);
            for (int s = 0; s < 16; s++) {
                uint32_t carry = B.p[s];
                for (int b = 0; b < nb && carry; b++) {
                    uint32_t x = h[s][b];
                    h[s][b] = x ^ carry;
                    carry &= x;
                }
            }
        }
        uint32_t cand[16], mx[kPolyBits];
        for (int s = 0; s < 16; s++) cand[s] = ~0u;
        for (int b = nb - 1; b >= 0; b--) {
            uint32_t one = 0;
            for (int s = 0; s < 16; s++) one |= cand[s] & h[s][b];
            for (int s = 0; s < 16; s++) cand[s] &= ~(one & ~h[s][b]);
            mx[b] = one;
        }
        uint32_t borrow = ~0u;  // mx - 1
        for (int b = 0; b < nb; b++) {
            uint32_t x = mx[b];
            mx[b] = x ^ borrow;
            borrow &= ~x;
        }
        for (int s = 0; s < 16; s++) {
            uint32_t eq = ~borrow;  // borrow out set <=> max was 0 (no valid child): D1 is irrelevant there
            for (int b = 0; b < nb; b++) eq &= ~(h[s][b] ^ mx[b]);
            S.p[s] = cand[s] & any_valid;
            D1.p[s] = eq;
        }
    }
}

template <int ALGO, bool POLY, class Env>
PMFS_HD void forward_walk(Env& env, const uint32_t* prog, int32_t len, const Code4* forced) {
    Planes16 T;
    zero16(T);
    int32_t pc = 0;
    while (pc < len) {
        const uint32_t w = env.word(prog, pc++);
        const uint32_t code = w & ((1u << kOpBits) - 1), id = w >> kOpBits;
        if (code == F_PUSH) {
            env.push(T);
            continue;
        }
        Planes16 D1;
        if (code == F_POLY) {
            if (POLY) {
                const uint32_t k = env.word(prog, pc++);
                poly_join<ALGO>(env, prog, pc, k, T, D1);
                pc += (int32_t)k;
            }
        } else {
            Planes16 A;
            uint32_t any;
            if (code == F_JOIN_POP) {
                env.pop(A);
            } else if (code == F_JOIN_IMM) {
                load_operand(env, env.word(prog, pc++), A);
            } else {  // F_JOIN_IMM2
                load_operand(env, env.word(prog, pc++), A);
                load_operand(env, env.word(prog, pc++), T);
            }
            if (ALGO == kAlgoSankoff) {
                Planes16 S;
                join_sets(A, T, S, any);
                join_d1(A, T, any, D1);
                T = S;
            } else {
                join_sets(A, T, T, any);
            }
        }
        finish_node<ALGO>(env, id, T, forced);
        if (ALGO == kAlgoSankoff) env.store_d(id, D1);
    }
}

// P/pv: state of the chunk root's parent (consensus planes and all-ones for the tree root).  colmask: real columns of
// this group.  Returns the mask of real columns whose Sankoff root has no finite cost and is not forced (an error).
template <int ALGO, class Env>
PMFS_HD uint32_t backward_walk(Env& env, const uint32_t* prog, int32_t len, Code4 P, uint32_t pv, const Code4* forced,
                               uint32_t colmask) {
    uint32_t err = 0;
    int32_t pc = 0;
    while (pc < len) {
        const uint32_t w = env.word(prog, pc++);
        const uint32_t code = w & ((1u << kOpBits) - 1), arg = w >> kOpBits;
        if (code == B_ENTER) {
            Planes16 S;
            env.load_x(arg, S);
            const uint32_t hit = mux16(S, P);  // parent's state is in this node's set
            const uint32_t nz = or16(S);
            if (ALGO == kAlgoFitch) {
                // a node with an empty set is skipped with its whole subtree (:101-103); below it every set is empty too
                const uint32_t m = nz & ~hit & colmask;
                if (env.any(m != 0)) {
                    Code4 l, G;
                    lowest(S, nz, l);
                    code_select(hit, P, l, G);
                    env.emit_internal(arg, m, G, P);
                    P = G;
                }
            } else {
                const uint32_t slow = pv & ~hit & colmask;  // includes invalid nodes (empty S never hits)
                if (env.any(slow != 0)) {
                    Planes16 D1;
                    env.load_d(arg, D1);
                    const uint32_t ind = mux16(D1, P);
                    Code4 l, G;
                    lowest(S, nz, l);
                    const uint32_t keep = hit | (ind & ~code_lt(l, P));  // d[p]==1: first index of {p} U S wins
                    code_select(keep, P, l, G);
                    const uint32_t nv = pv & nz;
                    env.emit_internal(arg, nv & code_ne(G, P) & colmask, G, P);
                    P = G;
                    pv = nv;
                }
            }
        } else if (code == B_LEAF) {
            Code4 c;
            transpose_nibbles(env.leaf(arg), c);
            // Fitch: parent & {c} ? parent : c == c.  Sankoff: S = {c}, D1 = {} -> c.  Either way the leaf keeps its code.
            uint32_t m = code_ne(c, P) & colmask;
            if (ALGO == kAlgoSankoff) m &= pv;
            if (env.any(m != 0)) env.emit_leaf(arg, m, c, P);
        } else if (code == B_XCHILD) {
            env.store_x(arg, P, pv);
        } else if (code == B_PUSHP) {
            env.pushp(P, pv);
        } else if (code == B_POPP) {
            env.popp(P, pv);
        } else {  // B_ENTER_ROOT: P holds the consensus
            Planes16 S;
            env.load_x(0, S);
            const uint32_t nz = or16(S);
            Code4 G;
            uint32_t nv;
            if (ALGO == kAlgoSankoff && forced) {
                G = *forced;  // defaultValue (:491-492)
                nv = ~0u;
            } else {
                lowest(S, nz, G);  // Fitch: lowest set bit (:104-116); Sankoff: first-index argmin (:493-507)
                nv = nz;
                if (ALGO == kAlgoSankoff) err |= ~nz & colmask;
            }
            const uint32_t m = nv & code_ne(G, P) & colmask;
            if (env.any(m != 0)) env.emit_internal(0, m, G, P);
            P = G;
            pv = nv;
        }
    }
    return err;
}

}  // namespace pmfs
