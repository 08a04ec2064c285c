/* TEST INFRASTRUCTURE ONLY (oracle).  Not part of the product: only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load the library built from this file.  The product (panman_b200/) never
 * links, loads or calls it, and has no CPU fallback.
 *
 * Plain-C restatement of the reference's per-column small-parsimony path, array based and non-recursive, so that it
 * runs where /root/reference does not exist (the GPU box) and on trees too deep for the reference's recursion.
 * It deliberately keeps the reference's own formulation (Fitch on 16-bit masks, Sankoff on 16 x int32 cost vectors with
 * SANKOFF_INF) -- NOT the two-mask form the CUDA kernels use -- so that it is an independent check of that rewrite.
 * Parity is PINNED: tests/test_oracle.py checks this file against oracle/_ref (the reference's own fitchSankoff.cpp
 * compiled in place) in this container, and against the golden fixtures in tests/golden/ generated from oracle/_ref.
 *
 * Each function cites the reference lines it follows (paths relative to /root/reference):
 *   code tables                  src/panman.cpp:46-81, :83-118
 *   Fitch forward                src/fitchSankoff.cpp:30-56
 *   Fitch backward               src/fitchSankoff.cpp:96-129
 *   Fitch assign                 src/fitchSankoff.cpp:131-171
 *   Sankoff forward              src/fitchSankoff.cpp:359-405
 *   Sankoff backward             src/fitchSankoff.cpp:487-531
 *   Sankoff assign               src/fitchSankoff.cpp:676-703
 *   per-column state build       src/panman.cpp:1397-1415 (Fitch), :1564-1596 (Sankoff)
 *   sort + <=6-run packing       src/panman.cpp:1435-1456; NucMut range ctor src/panman.hpp:86-128
 *
 * Tree convention (same as include/pmfs.h): node ids are the reference's creation order (pre-order; id 0 is the root
 * "node_1"), so parent[i] < i; children in Newick order as CSR; leaf_row[i] = MSA row of a leaf, -1 for internal
 * nodes, -2 for a leaf that has no row in the MSA.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define SANKOFF_INF 100000001 /* src/common.hpp:16 */

typedef struct {
    int32_t node;
    int32_t pos;
    uint8_t type; /* NS=0 ND=1 NI=2 (src/panman.hpp:26-41) */
    uint8_t code;
    uint8_t pad[2];
} pmo_mut;

typedef struct {
    int32_t nucPosition;
    int32_t nucGapPosition;
    int32_t primaryBlockId;
    int32_t secondaryBlockId;
    uint32_t mutInfo;
    uint32_t nucs;
} pmo_nucmut;

char pmo_nucleotide_from_code(int code) { /* src/panman.cpp:46-81 */
    switch (code) {
        case 1: return 'A';  case 2: return 'C';  case 4: return 'G';  case 8: return 'T';
        case 5: return 'R';  case 10: return 'Y'; case 6: return 'S';  case 9: return 'W';
        case 12: return 'K'; case 3: return 'M';  case 14: return 'B'; case 13: return 'D';
        case 11: return 'H'; case 7: return 'V';  case 15: return 'N';
        default: return '-';
    }
}

int pmo_code_from_nucleotide(char nuc) { /* src/panman.cpp:83-118 */
    switch (nuc) {
        case 'A': return 1;  case 'C': return 2;  case 'G': return 4;  case 'T': return 8;
        case 'R': return 5;  case 'Y': return 10; case 'S': return 6;  case 'W': return 9;
        case 'K': return 12; case 'M': return 3;  case 'B': return 14; case 'D': return 13;
        case 'H': return 11; case 'V': return 7;  case 'N': return 15;
        default: return 0;
    }
}

void pmo_free(void* p) { free(p); }

typedef struct {
    pmo_mut* v;
    int64_t n, cap;
} mutvec;

static void push(mutvec* m, int32_t node, int32_t pos, int type, char nuc) {
    if (m->n == m->cap) {
        m->cap = m->cap ? m->cap * 2 : 1024;
        m->v = (pmo_mut*)realloc(m->v, (size_t)m->cap * sizeof(pmo_mut));
    }
    pmo_mut* r = &m->v[m->n++];
    r->node = node;
    r->pos = pos;
    r->type = (uint8_t)type;
    r->code = (uint8_t)pmo_code_from_nucleotide(nuc); /* src/panman.cpp:1418: getCodeFromNucleotide(mutation char) */
    r->pad[0] = r->pad[1] = 0;
}

typedef struct {
    int32_t V;
    const int32_t *parent, *child_off, *child_idx, *leaf_row;
    const uint8_t* msa; /* [n_rows][n_cols] codes 0..15, one byte per cell */
    int64_t n_cols;
    const uint8_t* consensus; /* [n_cols] codes */
    const uint8_t* forced;    /* [n_cols] codes or NULL */
} problem;

/* One column of the Fitch branch.  st[] is the reference's `states` map as an array. */
static void fitch_column(const problem* P, int64_t col, int32_t* st, uint8_t* visited, mutvec* out) {
    int32_t V = P->V;
    /* forward (fitchSankoff.cpp:30-56): children before parents == descending pre-order id */
    for (int32_t v = V - 1; v >= 0; v--) {
        int32_t b = P->child_off[v], e = P->child_off[v + 1];
        if (b == e) {
            /* leaf: its mask (panman.cpp:1398-1406: gap and every unknown character -> bit 0), 0 if absent (:33-36) */
            st[v] = (P->leaf_row[v] >= 0) ? (1 << P->msa[(size_t)P->leaf_row[v] * P->n_cols + col]) : 0;
            continue;
        }
        if (P->parent[v] < 0 && P->forced) { /* :45-47 root overwritten with refState */
            st[v] = 1 << P->forced[col];
            continue;
        }
        int32_t orS = 0, andS = st[P->child_idx[b]];
        for (int32_t k = b; k < e; k++) {
            orS |= st[P->child_idx[k]];
            andS &= st[P->child_idx[k]];
        }
        st[v] = andS ? andS : orS;
    }
    /* backward (:96-129; the MSA path never passes defaultState) and assign (:131-171), both top-down */
    for (int32_t v = 0; v < V; v++) {
        int32_t p = P->parent[v];
        if (p >= 0 && !visited[p]) { visited[v] = 0; continue; } /* recursion never reached this node */
        if (st[v] == 0) { visited[v] = 0; continue; }            /* :101-103 return: subtree skipped */
        int32_t parentState = (p < 0) ? (1 << P->consensus[col]) : st[p];
        if (p < 0 || !(parentState & st[v])) {
            int32_t cur = 1; /* lowest set bit, :107-116 / :120-124 */
            while (!(st[v] & cur)) cur <<= 1;
            st[v] = cur;
        } else {
            st[v] = parentState;
        }
        visited[v] = 1;
        if (parentState != st[v]) { /* :139-167 */
            if (parentState == 1 || st[v] != 1) {
                int code = 0, cur = st[v];
                while (cur > 0) { cur >>= 1; code++; }
                code--;
                push(out, v, (int32_t)col, parentState == 1 ? 2 : 0, pmo_nucleotide_from_code(code));
            } else {
                push(out, v, (int32_t)col, 1, '-');
            }
        }
    }
}

/* One column of the Sankoff branch.  cost[] is the reference's `stateSets`, st[] its `states` (-1 = none). */
static int sankoff_column(const problem* P, int64_t col, int32_t* cost, int32_t* st, mutvec* out) {
    int32_t V = P->V;
    /* forward (:359-405) */
    for (int32_t v = V - 1; v >= 0; v--) {
        int32_t* c = cost + (size_t)v * 16;
        int32_t b = P->child_off[v], e = P->child_off[v + 1];
        if (b == e) {
            for (int i = 0; i < 16; i++) c[i] = SANKOFF_INF; /* blank if absent (:363-366) */
            if (P->leaf_row[v] >= 0) c[P->msa[(size_t)P->leaf_row[v] * P->n_cols + col]] = 0; /* panman.cpp:1565-1571 */
            continue;
        }
        int minExists = 0;
        for (int32_t k = b; k < e && !minExists; k++)
            for (int i = 0; i < 16; i++)
                if (cost[(size_t)P->child_idx[k] * 16 + i] < SANKOFF_INF) { minExists = 1; break; }
        if (!minExists) {
            for (int i = 0; i < 16; i++) c[i] = SANKOFF_INF;
            continue;
        }
        for (int i = 0; i < 16; i++) {
            int32_t acc = 0;
            for (int32_t k = b; k < e; k++) {
                const int32_t* cc = cost + (size_t)P->child_idx[k] * 16;
                int32_t minVal = SANKOFF_INF;
                for (int j = 0; j < 16; j++) {
                    int32_t x = (i != j) + cc[j];
                    if (x < minVal) minVal = x;
                }
                if (minVal < SANKOFF_INF) acc += minVal;
            }
            c[i] = acc;
        }
    }
    /* backward (:487-531) + assign (:676-703), top-down */
    for (int32_t v = 0; v < V; v++) {
        int32_t p = P->parent[v];
        const int32_t* c = cost + (size_t)v * 16;
        if (p < 0) {
            if (P->forced) {
                st[v] = P->forced[col]; /* defaultValue (panman.cpp:1573-1594) */
            } else {
                int32_t minVal = SANKOFF_INF, minPtr = -1;
                for (int i = 0; i < 16; i++)
                    if (c[i] < minVal) { minVal = c[i]; minPtr = i; }
                if (minPtr == -1) return -4; /* the reference asserts (:505) */
                st[v] = minPtr;
            }
        } else if (st[p] == -1) {
            st[v] = -1; /* :510-513 */
        } else {
            int32_t minPtr = -1, minVal = SANKOFF_INF; /* :515-527: strict <, first index wins */
            for (int i = 0; i < 16; i++) {
                int32_t x = (i != st[p]) + c[i];
                if (x < minVal) { minVal = x; minPtr = i; }
            }
            st[v] = minPtr;
        }
        if (st[v] == -1) continue; /* assign :680-682 (the whole subtree is -1) */
        int32_t parentState = (p < 0) ? P->consensus[col] : st[p];
        if (parentState != st[v]) {
            if (parentState == 0) push(out, v, (int32_t)col, 2, pmo_nucleotide_from_code(st[v]));
            else if (st[v] == 0) push(out, v, (int32_t)col, 1, '-');
            else push(out, v, (int32_t)col, 0, pmo_nucleotide_from_code(st[v]));
        }
    }
    return 0;
}

typedef struct {
    const problem* P;
    int algo;
    int64_t next; /* atomically advanced */
    mutvec out;
    int rc;
    pthread_mutex_t* mu;
    mutvec* global;
    int* grc;
    int64_t* gnext;
} worker;

static void* work(void* arg) {
    worker* w = (worker*)arg;
    const problem* P = w->P;
    int32_t* st = (int32_t*)malloc((size_t)P->V * sizeof(int32_t));
    uint8_t* vis = (uint8_t*)malloc((size_t)P->V);
    int32_t* cost = w->algo ? (int32_t*)malloc((size_t)P->V * 16 * sizeof(int32_t)) : NULL;
    mutvec local = {0, 0, 0};
    int rc = 0;
    for (;;) {
        int64_t begin = __atomic_fetch_add(w->gnext, 16, __ATOMIC_RELAXED);
        if (begin >= P->n_cols) break;
        int64_t end = begin + 16 < P->n_cols ? begin + 16 : P->n_cols;
        for (int64_t c = begin; c < end; c++) {
            if (w->algo == 0) fitch_column(P, c, st, vis, &local);
            else if (sankoff_column(P, c, cost, st, &local)) rc = -4;
        }
    }
    pthread_mutex_lock(w->mu);
    if (rc) *w->grc = rc;
    for (int64_t i = 0; i < local.n; i++) {
        mutvec* g = w->global;
        if (g->n == g->cap) {
            g->cap = g->cap ? g->cap * 2 : 1024;
            g->v = (pmo_mut*)realloc(g->v, (size_t)g->cap * sizeof(pmo_mut));
        }
        g->v[g->n++] = local.v[i];
    }
    pthread_mutex_unlock(w->mu);
    free(local.v);
    free(st);
    free(vis);
    free(cost);
    return NULL;
}

static int cmp_mut(const void* a, const void* b) {
    const pmo_mut *x = (const pmo_mut*)a, *y = (const pmo_mut*)b;
    if (x->node != y->node) return x->node < y->node ? -1 : 1;
    if (x->pos != y->pos) return x->pos < y->pos ? -1 : 1;
    return 0;
}

/* The column loop (src/panman.cpp:1371-1425 Fitch / :1558-1603 Sankoff) over codes.  Output sorted by (node,pos),
 * which is the per-node std::sort order of :1437.  Returns 0 or -4 (Sankoff root has no finite cost and is not forced). */
int pmo_run(int algo, int32_t n_nodes, const int32_t* parent, const int32_t* child_off, const int32_t* child_idx,
            const int32_t* leaf_row, const uint8_t* msa_codes, int64_t n_cols, const uint8_t* consensus_code,
            const uint8_t* root_forced_code, int n_threads, pmo_mut** out, int64_t* n_out) {
    problem P = {n_nodes, parent, child_off, child_idx, leaf_row, msa_codes, n_cols, consensus_code, root_forced_code};
    if (n_threads < 1) n_threads = 1;
    pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
    mutvec global = {0, 0, 0};
    int grc = 0;
    int64_t gnext = 0;
    worker* ws = (worker*)calloc((size_t)n_threads, sizeof(worker));
    pthread_t* th = (pthread_t*)calloc((size_t)n_threads, sizeof(pthread_t));
    for (int t = 0; t < n_threads; t++) {
        ws[t].P = &P;
        ws[t].algo = algo;
        ws[t].mu = &mu;
        ws[t].global = &global;
        ws[t].grc = &grc;
        ws[t].gnext = &gnext;
        pthread_create(&th[t], NULL, work, &ws[t]);
    }
    for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
    free(ws);
    free(th);
    if (grc) {
        free(global.v);
        *out = NULL;
        *n_out = 0;
        return grc;
    }
    qsort(global.v, (size_t)global.n, sizeof(pmo_mut), cmp_mut);
    if (!global.v) global.v = (pmo_mut*)malloc(sizeof(pmo_mut));
    *out = global.v;
    *n_out = global.n;
    return 0;
}

/* Greedy <=6-run packing of records already sorted by (node,pos): src/panman.cpp:1439-1455 + src/panman.hpp:86-128. */
int pmo_pack(const pmo_mut* muts, int64_t n_muts, int32_t n_nodes, int64_t** nuc_off, pmo_nucmut** nuc_out, int64_t* n_nuc) {
    int64_t* off = (int64_t*)calloc((size_t)n_nodes + 1, sizeof(int64_t));
    pmo_nucmut* packed = (pmo_nucmut*)malloc((size_t)(n_muts > 0 ? n_muts : 1) * sizeof(pmo_nucmut));
    int64_t np = 0, a = 0;
    for (int32_t node = 0; node < n_nodes; node++) {
        off[node] = np;
        int64_t b = a;
        while (b < n_muts && muts[b].node == node) b++;
        int64_t start = a;
        for (int64_t i = a; i <= b && b > a; i++) {
            int brk = (i == b) || (i > start && (i - start == 6 || muts[i].pos != muts[i - 1].pos + 1 ||
                                                 muts[i].type != muts[i - 1].type));
            if (brk) {
                pmo_nucmut m;
                m.primaryBlockId = 0;
                m.secondaryBlockId = -1;
                m.mutInfo = (uint8_t)(((i - start) << 4) + muts[start].type);
                m.nucPosition = muts[start].pos;
                m.nucGapPosition = -1;
                m.nucs = 0;
                for (int64_t k = start; k < i; k++) m.nucs += ((uint32_t)muts[k].code << (4 * (5 - (k - start))));
                packed[np++] = m;
                start = i;
            }
        }
        a = b;
    }
    off[n_nodes] = np;
    *nuc_off = off;
    *nuc_out = packed;
    *n_nuc = np;
    return 0;
}
