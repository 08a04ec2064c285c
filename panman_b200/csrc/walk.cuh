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

PMFS_HD uint32_t or16(const Planes16& s) {  // tree shaped: short dependency chain
    const uint32_t a = s.p[0] | s.p[1] | s.p[2], b = s.p[3] | s.p[4] | s.p[5], c = s.p[6] | s.p[7] | s.p[8];
    const uint32_t d = s.p[9] | s.p[10] | s.p[11], e = s.p[12] | s.p[13] | s.p[14];
    return (a | b | c) | (d | e | s.p[15]);
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
    uint32_t t[4] = {0, 0, 0, 0};  // four short OR chains instead of one of length 16
#pragma unroll
    for (int i = 0; i < 16; i++) t[i & 3] |= a.p[i] & b.p[i];
    const uint32_t any = (t[0] | t[1]) | (t[2] | t[3]);
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
// One thread walks NG column groups at once (NG x 32 columns): the per-op bookkeeping (program fetch, dispatch, operand
// hand-over, addressing) is paid once per NG groups, and the NG independent dependency chains interleave in the
// pipeline.  On the B200 this code is bound by integer-ALU issue, so instructions per column are what counts.
//
// Env concept (device: walk_kernels.cuh DevEnv; host emulation: tests/host_emul.cpp), NG = Env::kGroups
//   uint32_t word(const uint32_t* prog, int32_t pc)            program word, same for every thread
//   void     leaf(uint32_t row, uint4 (&w)[NG])                this thread's 32 nibbles per group of MSA row `row`
//   void     load_x(uint32_t id, Planes16 (&s)[NG])            set S[id] written by an earlier work item (prefetched)
//   void     load_s / store_s(uint32_t id, Planes16 (&s)[NG])  set array S (Fitch: the only one), direct access
//   void     load_d / store_d                                  Sankoff D1 array
//   void     push / pop(Planes16 (&s)[NG])                     forward stack
//   void     pushp / popp(Code4 (&c)[NG], uint32_t (&pv)[NG])  backward stack
//   void     store_x(uint32_t chunk, const Code4 (&c)[NG], const uint32_t (&pv)[NG])   parent state for a child chunk
//   bool     any(bool)                                         warp-wide OR (host: identity)
//   void     emit_internal(uint32_t id, const uint32_t (&m)[NG], const Code4 (&g)[NG], const Code4 (&p)[NG])
//   void     emit_leaf(uint32_t row, ...same...)
// ---------------------------------------------------------------------------------------------------------------------

template <class Env>
PMFS_HD void load_operand(Env& env, uint32_t o, Planes16 (&s)[Env::kGroups]) {
    constexpr int NG = Env::kGroups;
    if (o & kOperandSet) {
        if (o & kOperandExt) env.load_x(o & kOperandIdMask, s);  // written by an earlier work item: comes through the ring
        else env.load_s(o & kOperandIdMask, s);                  // written by this thread earlier in this work item
    } else if (o == kAbsentLeaf) {
#pragma unroll
        for (int j = 0; j < NG; j++) zero16(s[j]);
    } else {
        uint4 w[NG];
        env.leaf(o, w);
#pragma unroll
        for (int j = 0; j < NG; j++) {
            Code4 c;
            transpose_nibbles(w[j], c);
            onehot(c, s[j]);
        }
    }
}

template <int ALGO, class Env>
PMFS_HD void finish_node(Env& env, uint32_t id, Planes16 (&S)[Env::kGroups], const Code4* forced) {
    if (ALGO == kAlgoFitch && id == 0 && forced) {  // refState overwrites the root (:45-47)
#pragma unroll
        for (int j = 0; j < Env::kGroups; j++) onehot(forced[j], S[j]);
    }
    env.store_s(id, S);
}

// Any-arity join (unary nodes and polytomies); children are re-read from memory, so nothing is live across them.
// mode 0 (F_POLY): all k children of node `id`.  mode 1 (F_PART): fold the k children of one part of a split polytomy and
// store the partial result in the extra set rows from `id` on; S / D1 are not touched.  mode 2 (F_COMBINE): the k
// operands are parts (first extra row of each); combine their partial results into S / D1 of the node.
// (The mode is a template parameter: with a run-time mode the three variants' live ranges add up -- +30 registers on the
// kernels of every tree with a polytomy, split or not.)
template <int ALGO, int MODE, class Env>
PMFS_HD void poly_join(Env& env, const uint32_t* prog, int32_t pc, uint32_t k, uint32_t id, Planes16 (&S)[Env::kGroups],
                       Planes16 (&D1)[Env::kGroups]) {
    constexpr int NG = Env::kGroups;
    constexpr int mode = MODE;
    if (ALGO == kAlgoFitch) {
        // S = AND of the children's sets if that is not empty, else their OR (fitchSankoff.cpp:39-55): (AND, OR) of a
        // subset of the children is all a part has to hand on
        Planes16 A[NG], O[NG];
#pragma unroll
        for (int j = 0; j < NG; j++)
#pragma unroll
            for (int i = 0; i < 16; i++) { A[j].p[i] = ~0u; O[j].p[i] = 0; }
        for (uint32_t c = 0; c < k; c++) {
            const uint32_t o = env.word(prog, pc + (int32_t)c);
            Planes16 B[NG];
            if (mode == 2) {
                env.load_x(o & kOperandIdMask, B);
#pragma unroll
                for (int j = 0; j < NG; j++)
#pragma unroll
                    for (int i = 0; i < 16; i++) A[j].p[i] &= B[j].p[i];
                env.load_x((o & kOperandIdMask) + 1, B);
#pragma unroll
                for (int j = 0; j < NG; j++)
#pragma unroll
                    for (int i = 0; i < 16; i++) O[j].p[i] |= B[j].p[i];
                continue;
            }
            load_operand(env, o, B);
#pragma unroll
            for (int j = 0; j < NG; j++)
#pragma unroll
                for (int i = 0; i < 16; i++) { A[j].p[i] &= B[j].p[i]; O[j].p[i] |= B[j].p[i]; }
        }
        if constexpr (MODE == 1) {
            env.store_s(id, A);
            env.store_s(id + 1, O);
        } else {
#pragma unroll
            for (int j = 0; j < NG; j++) {
                const uint32_t any = or16(A[j]);
#pragma unroll
                for (int i = 0; i < 16; i++) S[j].p[i] = A[j].p[i] | (O[j].p[i] & ~any);
            }
        }
    } else {
        // h[s] = number of valid children whose S contains s, as bit-sliced counters; S = argmax h, D1 = {h = max-1}.
        // Children are counted three at a time into a 2-bit counter per state that lives in registers (statically
        // indexed: everything over the 16 states is unrolled, 16 independent chains, no control flow) -- three
        // increments cannot overflow it.  Arity <= 3 needs nothing else.  For larger arities the register counters are
        // added into full-width counters in a per-thread array after every third child (only the states some column of
        // the warp has seen; the carry ripples two bits on average), and read back from there at the end.
        // A part counts at most 63 children: its 16 counters are 6 bits, handed on as 6 set rows (row b = bit b of all
        // counters); the combine adds the parts' counters up.  (A valid child is in at least one counter, so "some child
        // is valid" is handed on with them.)
        const uint32_t k_all = mode == 2 ? k * 63u : k;  // what the counters must be able to hold
        int nb = 1;
        while ((k_all >> nb) != 0) nb++;
        const int nhi = nb > 2 ? nb - 2 : 0;
        uint32_t lo0[NG][16], lo1[NG][16], full[NG][16][kPolyBits];
        uint32_t any_valid[NG];
#pragma unroll
        for (int j = 0; j < NG; j++) {
            any_valid[j] = 0;
#pragma unroll
            for (int s_ = 0; s_ < 16; s_++) lo0[j][s_] = lo1[j][s_] = 0;
            if (nhi > 0) {
#pragma unroll
                for (int s_ = 0; s_ < 16; s_++)
                    for (int b_ = 0; b_ < nb; b_++) full[j][s_][b_] = 0;
            }
        }
        uint32_t pending = 0;  // children counted in lo0/lo1 since the last flush
        for (uint32_t c = 0; c < k; c++) {
            if (mode == 2) {  // full += the 6-bit counters of part c (ripple-carry over the bit rows, then into the high bits)
                const uint32_t row = env.word(prog, pc + (int32_t)c) & kOperandIdMask;
                uint32_t cy[NG][16];
#pragma unroll
                for (int j = 0; j < NG; j++)
#pragma unroll
                    for (int s_ = 0; s_ < 16; s_++) cy[j][s_] = 0;
                for (int b_ = 0; b_ < kPartRowsSankoff; b_++) {
                    Planes16 R[NG];
                    env.load_x(row + (uint32_t)b_, R);
#pragma unroll
                    for (int j = 0; j < NG; j++) {
                        any_valid[j] |= or16(R[j]);
#pragma unroll
                        for (int s_ = 0; s_ < 16; s_++) {
                            const uint32_t x = full[j][s_][b_], r = R[j].p[s_], c_in = cy[j][s_];
                            full[j][s_][b_] = x ^ r ^ c_in;
                            cy[j][s_] = (x & r) | (c_in & (x ^ r));
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < NG; j++)
#pragma unroll
                    for (int s_ = 0; s_ < 16; s_++) {
                        uint32_t carry = cy[j][s_];
                        for (int b_ = kPartRowsSankoff; b_ < nb && env.any(carry != 0); b_++) {
                            const uint32_t x = full[j][s_][b_];
                            full[j][s_][b_] = x ^ carry;
                            carry &= x;
                        }
                    }
                continue;
            }
            Planes16 B[NG];
            load_operand(env, env.word(prog, pc + (int32_t)c), B);
#pragma unroll
            for (int j = 0; j < NG; j++) {
                any_valid[j] |= or16(B[j]);
#pragma unroll
                for (int s_ = 0; s_ < 16; s_++) {
                    const uint32_t x = lo0[j][s_], carry = B[j].p[s_] & x;
                    lo0[j][s_] = x ^ B[j].p[s_];
                    lo1[j][s_] ^= carry;
                }
            }
            pending++;
            if (nhi > 0 && (pending == 3 || c + 1 == k)) {
                pending = 0;
#pragma unroll
                for (int j = 0; j < NG; j++)
#pragma unroll
                    for (int s_ = 0; s_ < 16; s_++) {
                        const uint32_t a0 = lo0[j][s_], a1 = lo1[j][s_];
                        if (!env.any((a0 | a1) != 0)) continue;
                        uint32_t x = full[j][s_][0];
                        full[j][s_][0] = x ^ a0;
                        uint32_t carry = x & a0;
                        x = full[j][s_][1];
                        full[j][s_][1] = x ^ a1 ^ carry;
                        carry = (x & a1) | (carry & (x ^ a1));
                        for (int b_ = 2; b_ < nb && env.any(carry != 0); b_++) {
                            x = full[j][s_][b_];
                            full[j][s_][b_] = x ^ carry;
                            carry &= x;
                        }
                        lo0[j][s_] = lo1[j][s_] = 0;
                    }
            }
        }
        if (nhi > 0) {  // the two low bits back into registers: the code below reads (lo0, lo1, full[2..])
#pragma unroll
            for (int j = 0; j < NG; j++)
#pragma unroll
                for (int s_ = 0; s_ < 16; s_++) {
                    lo0[j][s_] = full[j][s_][0];
                    lo1[j][s_] = full[j][s_][1];
                }
        }
        if constexpr (MODE == 1) {
            for (int b_ = 0; b_ < kPartRowsSankoff; b_++) {
                Planes16 R[NG];
#pragma unroll
                for (int j = 0; j < NG; j++)
#pragma unroll
                    for (int s_ = 0; s_ < 16; s_++)
                        R[j].p[s_] = b_ == 0 ? lo0[j][s_] : b_ == 1 ? lo1[j][s_] : (b_ < nb ? full[j][s_][b_] : 0u);
                env.store_s(id + (uint32_t)b_, R);
            }
        } else {
#pragma unroll
        for (int j = 0; j < NG; j++) {
            // argmax over the 16 counters, most significant bit first: `cand` keeps the states that can still be maximal,
            // mx collects the maximum's bits.  The two register-resident bits are straight-line code; the loop over the
            // high bits does not run for arity <= 3.
            uint32_t cand[16], mxhi[kPolyBits - 2], mx0, mx1;
#pragma unroll
            for (int s_ = 0; s_ < 16; s_++) cand[s_] = ~0u;
            for (int b_ = nhi - 1; b_ >= 0; b_--) {
                uint32_t one = 0;
#pragma unroll
                for (int s_ = 0; s_ < 16; s_++) one |= cand[s_] & full[j][s_][b_ + 2];
#pragma unroll
                for (int s_ = 0; s_ < 16; s_++) cand[s_] &= ~(one & ~full[j][s_][b_ + 2]);
                mxhi[b_] = one;
            }
            {
                uint32_t one = 0;
#pragma unroll
                for (int s_ = 0; s_ < 16; s_++) one |= cand[s_] & lo1[j][s_];
#pragma unroll
                for (int s_ = 0; s_ < 16; s_++) cand[s_] &= ~(one & ~lo1[j][s_]);
                mx1 = one;
                one = 0;
#pragma unroll
                for (int s_ = 0; s_ < 16; s_++) one |= cand[s_] & lo0[j][s_];
#pragma unroll
                for (int s_ = 0; s_ < 16; s_++) cand[s_] &= ~(one & ~lo0[j][s_]);
                mx0 = one;
            }
            // mx - 1 (borrow chain from bit 0); a borrow out of the top bit <=> max was 0 (no valid child): D1 is
            // irrelevant there
            uint32_t borrow = ~mx0;
            mx0 = ~mx0;
            {
                const uint32_t x = mx1;
                mx1 = x ^ borrow;
                borrow &= ~x;
            }
            for (int b_ = 0; b_ < nhi; b_++) {
                const uint32_t x = mxhi[b_];
                mxhi[b_] = x ^ borrow;
                borrow &= ~x;
            }
            uint32_t eq[16];
#pragma unroll
            for (int s_ = 0; s_ < 16; s_++) eq[s_] = ~borrow & ~(lo0[j][s_] ^ mx0) & ~(lo1[j][s_] ^ mx1);
            for (int b_ = 0; b_ < nhi; b_++) {
#pragma unroll
                for (int s_ = 0; s_ < 16; s_++) eq[s_] &= ~(full[j][s_][b_ + 2] ^ mxhi[b_]);
            }
#pragma unroll
            for (int s_ = 0; s_ < 16; s_++) {
                S[j].p[s_] = cand[s_] & any_valid[j];
                D1[j].p[s_] = eq[s_];
            }
        }
        }
    }
}

// forced: NG code planes (Fitch refState) or nullptr
template <int ALGO, bool POLY, class Env>
PMFS_HD void forward_walk(Env& env, const uint32_t* prog, int32_t len, const Code4* forced) {
    constexpr int NG = Env::kGroups;
    Planes16 T[NG];
#pragma unroll
    for (int j = 0; j < NG; j++) zero16(T[j]);
    // Program words are fetched three ahead of their use (an op has at most two operand words): the load of a word --
    // shared memory for staged programs, L1/L2 for long ones -- is then never waited for.  (Reading up to three words past
    // the end of a program is harmless: programs are padded, see plan.cpp.)
    int32_t pc = 0;
    uint32_t q0 = env.word(prog, 0), q1 = env.word(prog, 1), q2 = env.word(prog, 2);
    auto take = [&]() {
        const uint32_t r = q0;
        q0 = q1;
        q1 = q2;
        q2 = env.word(prog, pc + 3);
        pc++;
        return r;
    };
    while (pc < len) {
        const uint32_t w = take();
        const uint32_t code = w & ((1u << kOpBits) - 1), id = w >> kOpBits;
        if (code == F_PUSH) {
            env.push(T);
            continue;
        }
        Planes16 D1[NG];
        if (code >= F_POLY) {  // F_POLY, F_PART, F_COMBINE
            if (POLY) {
                const uint32_t k = take();
                if (code == F_POLY) poly_join<ALGO, 0>(env, prog, pc, k, id, T, D1);
                else if (code == F_PART) poly_join<ALGO, 1>(env, prog, pc, k, id, T, D1);
                else poly_join<ALGO, 2>(env, prog, pc, k, id, T, D1);
                pc += (int32_t)k;
                q0 = env.word(prog, pc); q1 = env.word(prog, pc + 1); q2 = env.word(prog, pc + 2);
                if (code == F_PART) continue;  // a part stores its partial result itself; no node is finished here
            }
        } else {
            Planes16 A[NG];
            if (code == F_JOIN_POP) {
                env.pop(A);
            } else if (code == F_JOIN_IMM) {
                load_operand(env, take(), A);
            } else {  // F_JOIN_IMM2
                const uint32_t o1 = take(), o2 = take();
#ifdef PMFS_DBG_NO_CHERRY
                if (false) {
#else
                if (!((o1 | o2) & kOperandSet) && o1 != kAbsentLeaf && o2 != kAbsentLeaf) {
#endif
                    // a cherry (two leaves, a third of the internal nodes of a random binary tree): two one-element sets
                    // meet only if equal, and then their union is the same set -- S = {a} | {b} without a join;
                    // Sankoff's D1 is everything else where a != b and empty where a == b
                    uint4 wa[NG], wb[NG];
                    env.leaf(o1, wa);
                    env.leaf(o2, wb);
#pragma unroll
                    for (int j = 0; j < NG; j++) {
                        Code4 a, b;
                        transpose_nibbles(wa[j], a);
                        transpose_nibbles(wb[j], b);
                        uint32_t la[4], ha[4], lb[4], hb[4];
                        la[0] = ~a.b[1] & ~a.b[0]; la[1] = ~a.b[1] & a.b[0]; la[2] = a.b[1] & ~a.b[0]; la[3] = a.b[1] & a.b[0];
                        ha[0] = ~a.b[3] & ~a.b[2]; ha[1] = ~a.b[3] & a.b[2]; ha[2] = a.b[3] & ~a.b[2]; ha[3] = a.b[3] & a.b[2];
                        lb[0] = ~b.b[1] & ~b.b[0]; lb[1] = ~b.b[1] & b.b[0]; lb[2] = b.b[1] & ~b.b[0]; lb[3] = b.b[1] & b.b[0];
                        hb[0] = ~b.b[3] & ~b.b[2]; hb[1] = ~b.b[3] & b.b[2]; hb[2] = b.b[3] & ~b.b[2]; hb[3] = b.b[3] & b.b[2];
#pragma unroll
                        for (int s_ = 0; s_ < 16; s_++) T[j].p[s_] = (ha[s_ >> 2] & la[s_ & 3]) | (hb[s_ >> 2] & lb[s_ & 3]);
                        if (ALGO == kAlgoSankoff) {
                            const uint32_t ne = code_ne(a, b);
#pragma unroll
                            for (int s_ = 0; s_ < 16; s_++) D1[j].p[s_] = ~T[j].p[s_] & ne;
                        }
                    }
                    finish_node<ALGO>(env, id, T, forced);
                    if (ALGO == kAlgoSankoff) env.store_d(id, D1);
                    continue;
                }
                load_operand(env, o1, A);
                load_operand(env, o2, T);
            }
#pragma unroll
            for (int j = 0; j < NG; j++) {
                uint32_t any;
                if (ALGO == kAlgoSankoff) {
                    Planes16 S;
                    join_sets(A[j], T[j], S, any);
                    join_d1(A[j], T[j], any, D1[j]);
                    T[j] = S;
                } else {
                    join_sets(A[j], T[j], T[j], any);
                }
            }
        }
        finish_node<ALGO>(env, id, T, forced);
        if (ALGO == kAlgoSankoff) env.store_d(id, D1);
    }
}

// P/pv: state of the chunk root's parent (consensus planes and all-ones for the tree root), per group.  colmask: real
// columns of each group.  Returns the mask of real columns whose Sankoff root has no finite cost and is not forced.
template <int ALGO, class Env>
PMFS_HD uint32_t backward_walk(Env& env, const uint32_t* prog, int32_t len, Code4 (&P)[Env::kGroups],
                               uint32_t (&pv)[Env::kGroups], const Code4* forced, const uint32_t (&colmask)[Env::kGroups]) {
    constexpr int NG = Env::kGroups;
    uint32_t err = 0;
    int32_t pc = 0;
    uint32_t w_next = env.word(prog, 0);  // every op is one word: fetched one op ahead
    while (pc < len) {
        const uint32_t w = w_next;
        w_next = env.word(prog, ++pc);
        const uint32_t code = w & ((1u << kOpBits) - 1), arg = w >> kOpBits;
        if (code == B_ENTER) {
            Planes16 S[NG];
            env.load_x(arg, S);
            uint32_t hit[NG], nz[NG], slow = 0;
#pragma unroll
            for (int j = 0; j < NG; j++) {
                hit[j] = mux16(S[j], P[j]);  // parent's state is in this node's set
                nz[j] = or16(S[j]);
                // Fitch: a node with an empty set is skipped with its whole subtree (:101-103); below it every set is
                // empty too.  Sankoff: includes invalid nodes (an empty S never hits).
                slow |= (ALGO == kAlgoFitch ? nz[j] : pv[j]) & ~hit[j] & colmask[j];
            }
            if (env.any(slow != 0)) {
                Planes16 D1[NG];
                if (ALGO == kAlgoSankoff) env.load_d(arg, D1);
                Code4 G[NG];
                uint32_t m[NG];
#pragma unroll
                for (int j = 0; j < NG; j++) {
                    Code4 l;
                    lowest(S[j], nz[j], l);
                    if (ALGO == kAlgoFitch) {
                        code_select(hit[j], P[j], l, G[j]);
                        m[j] = nz[j] & ~hit[j] & colmask[j];
                    } else {
                        const uint32_t ind = mux16(D1[j], P[j]);
                        const uint32_t keep = hit[j] | (ind & ~code_lt(l, P[j]));  // d[p]==1: first index of {p} U S wins
                        code_select(keep, P[j], l, G[j]);
                        const uint32_t nv = pv[j] & nz[j];
                        m[j] = nv & code_ne(G[j], P[j]) & colmask[j];
                        pv[j] = nv;
                    }
                }
                env.emit_internal(arg, m, G, P);
#pragma unroll
                for (int j = 0; j < NG; j++) P[j] = G[j];
            }
        } else if (code == B_LEAF) {
            uint4 wv[NG];
            env.leaf(arg, wv);
            Code4 c[NG];
            uint32_t m[NG], anym = 0;
#pragma unroll
            for (int j = 0; j < NG; j++) {
                transpose_nibbles(wv[j], c[j]);
                // Fitch: parent & {c} ? parent : c == c.  Sankoff: S = {c}, D1 = {} -> c.  The leaf keeps its code.
                m[j] = code_ne(c[j], P[j]) & colmask[j];
                if (ALGO == kAlgoSankoff) m[j] &= pv[j];
                anym |= m[j];
            }
            if (env.any(anym != 0)) env.emit_leaf(arg, m, c, P);
        } else if (code == B_XCHILD) {
            env.store_x(arg, P, pv);
        } else if (code == B_PUSHP) {
            env.pushp(P, pv);
        } else if (code == B_POPP) {
            env.popp(P, pv);
        } else {  // B_ENTER_ROOT: P holds the consensus
            Planes16 S[NG];
            env.load_x(0, S);
            Code4 G[NG];
            uint32_t m[NG], anym = 0;
#pragma unroll
            for (int j = 0; j < NG; j++) {
                const uint32_t nz = or16(S[j]);
                uint32_t nv;
                if (ALGO == kAlgoSankoff && forced) {
                    G[j] = forced[j];  // defaultValue (:491-492)
                    nv = ~0u;
                } else {
                    lowest(S[j], nz, G[j]);  // Fitch: lowest set bit (:104-116); Sankoff: first-index argmin (:493-507)
                    nv = nz;
                    if (ALGO == kAlgoSankoff) err |= ~nz & colmask[j];
                }
                m[j] = nv & code_ne(G[j], P[j]) & colmask[j];
                anym |= m[j];
                pv[j] = nv;
            }
            if (env.any(anym != 0)) env.emit_internal(0, m, G, P);
#pragma unroll
            for (int j = 0; j < NG; j++) P[j] = G[j];
        }
    }
    return err;
}

}  // namespace pmfs
