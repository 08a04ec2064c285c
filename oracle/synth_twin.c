/* TEST INFRASTRUCTURE ONLY: CPU twin of libpmfs's device-side synthetic alignment generator
 * (panman_b200/csrc/aux_kernels.cuh: synth_root / synth_branch / synth_leaf, driven by pmfs_synth_msa in pmfs.cu).
 * The alignment is a pure function of (seed, node, column): simulated evolution down the tree, SURVEY §8(d).  The twin
 * lets bench.py's reference arm (the reference's CPU path on the host cores) produce its input without a GPU and
 * without loading the product library, and lets the tests check the device generator cell for cell.
 * It restates the generator, not the reference: nothing here comes from /root/reference. */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static uint64_t mix64(uint64_t x) { /* splitmix64 finaliser */
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
    x ^= x >> 27; x *= 0x94d049bb133111ebull;
    x ^= x >> 31;
    return x;
}
static uint32_t hash3(uint64_t seed, uint64_t a, uint64_t b) {
    return (uint32_t)(mix64(seed ^ mix64(a * 0x9e3779b97f4a7c15ull + b)) >> 32);
}

typedef struct {
    uint64_t seed;
    float p_sub, p_del, p_ins, p_n_block, p_amb, accessory_frac;
} pmo_synth_spec; /* = pmfs_synth_spec, include/pmfs.h */

typedef struct {
    uint64_t seed;
    uint32_t t_sub, t_del, t_ins, t_nblock, t_amb, t_acc;
    int32_t n_nodes, n_rows;
    const int32_t *parent, *leaf_row;
    const int64_t* cols;
    int64_t n_sel, next;
    uint8_t* out; /* [n_rows][n_sel] codes */
    pthread_mutex_t mu;
} Job;

static uint32_t thr(float p) {
    double x = (double)p * 4294967296.0;
    return x >= 4294967295.0 ? 0xffffffffu : (x <= 0 ? 0u : (uint32_t)x);
}

static const uint32_t base4[4] = {1, 2, 4, 8};

static uint32_t synth_root(const Job* s, int64_t col) {
    const uint32_t r = hash3(s->seed, 0x1000000000ull, (uint64_t)col), u = r >> 8;
    if (hash3(s->seed, 0x2000000000ull, (uint64_t)col) < s->t_acc) return 0;
    return u < 5033165u ? 1 : (u < 8053064u ? 2 : (u < 11408507u ? 4 : 8));
}
static uint32_t synth_branch(const Job* s, uint32_t st, int32_t node, int64_t col) {
    const uint32_t h = hash3(s->seed, (uint64_t)node, (uint64_t)col);
    if (st != 0) {
        if (h < s->t_sub) {
            const uint32_t nb = base4[(h >> 3) & 3];
            st = nb == st ? base4[((h >> 3) + 1) & 3] : nb;
        } else if (h - s->t_sub < s->t_del) {
            st = 0;
        }
    } else if (h < s->t_ins) {
        st = base4[(h >> 5) & 3];
    }
    return st;
}
static uint32_t synth_leaf(const Job* s, uint32_t st, int32_t row, int64_t col) {
    if (hash3(s->seed, 0x3000000000ull + (uint64_t)row, (uint64_t)(col >> 5)) < s->t_nblock) st = 15;
    else {
        const uint32_t h = hash3(s->seed, 0x4000000000ull + (uint64_t)row, (uint64_t)col);
        if (h < s->t_amb) {
            static const uint32_t amb[10] = {3, 5, 6, 7, 9, 10, 11, 12, 13, 14};
            st = amb[(h >> 7) % 10];
        }
    }
    if (st == 0 && (uint32_t)(hash3(s->seed, 0x5000000000ull, (uint64_t)col) % (uint32_t)s->n_rows) == (uint32_t)row)
        st = base4[hash3(s->seed, 0x1000000000ull, (uint64_t)col) & 3];
    return st;
}

static void* worker(void* p) {
    Job* s = (Job*)p;
    uint8_t* state = (uint8_t*)malloc((size_t)s->n_nodes);
    for (;;) {
        pthread_mutex_lock(&s->mu);
        const int64_t k = s->next++;
        pthread_mutex_unlock(&s->mu);
        if (k >= s->n_sel) break;
        const int64_t col = s->cols[k];
        /* ids are pre-order: a parent precedes its children */
        for (int32_t v = 0; v < s->n_nodes; v++) {
            uint32_t st = v == 0 ? synth_root(s, col) : synth_branch(s, state[s->parent[v]], v, col);
            const int32_t row = s->leaf_row[v];
            if (row >= 0) {
                st = synth_leaf(s, st, row, col);
                s->out[(size_t)row * (size_t)s->n_sel + (size_t)k] = (uint8_t)st;
            }
            state[v] = (uint8_t)st;
        }
    }
    free(state);
    return NULL;
}

/* codes of the selected global columns cols[0..n_sel) for every MSA row: out[row * n_sel + k] */
int pmo_synth(int32_t n_nodes, const int32_t* parent, const int32_t* leaf_row, int32_t n_rows, const pmo_synth_spec* spec,
              const int64_t* cols, int64_t n_sel, int n_threads, uint8_t* out) {
    if (n_nodes < 1 || !parent || !leaf_row || !spec || !cols || !out || n_rows < 1) return -1;
    Job s;
    memset(&s, 0, sizeof(s));
    s.seed = spec->seed;
    s.t_sub = thr(spec->p_sub); s.t_del = thr(spec->p_del); s.t_ins = thr(spec->p_ins);
    s.t_nblock = thr(spec->p_n_block); s.t_amb = thr(spec->p_amb); s.t_acc = thr(spec->accessory_frac);
    s.n_nodes = n_nodes; s.n_rows = n_rows; s.parent = parent; s.leaf_row = leaf_row;
    s.cols = cols; s.n_sel = n_sel; s.out = out;
    pthread_mutex_init(&s.mu, NULL);
    memset(out, 0, (size_t)n_rows * (size_t)n_sel);
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    pthread_t th[256];
    for (int t = 0; t < n_threads; t++) pthread_create(&th[t], NULL, worker, &s);
    for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
    pthread_mutex_destroy(&s.mu);
    return 0;
}
