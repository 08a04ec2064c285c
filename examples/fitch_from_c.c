/* The C ABI from plain C: a five-node tree, four alignment columns, Fitch, records printed one per line.
 *   gcc -std=c99 -Iinclude examples/fitch_from_c.c -Lpanman_b200 -lpmfs -Wl,-rpath,$PWD/panman_b200 -o /tmp/fitch_from_c
 * Needs a CUDA device: without one pmfs_create fails with a message (the library has no CPU path) and this program
 * prints it and exits 2 -- which is what tests/test_abi_cpu.py checks on the CPU box.
 *
 * Tree ((t0,t1),t2): nodes in pre-order 0 = root, 1 = (t0,t1), 2 = t0, 3 = t1, 4 = t2; MSA rows 0..2 = t0..t2.
 * Codes are the reference's 4-bit nucleotide codes (src/panman.cpp:83-118): A 1, C 2, G 4, T 8, N 15, gap 0; two
 * columns per byte, even column in the low nibble. */
#include <stdio.h>
#include <stdlib.h>

#include "pmfs.h"

int main(void) {
    static const int32_t parent[5] = {-1, 0, 1, 1, 0};
    static const int32_t child_off[6] = {0, 2, 4, 4, 4, 4};
    static const int32_t child_idx[4] = {1, 4, 2, 3};
    static const int32_t leaf_row[5] = {-1, -1, 0, 1, 2};
    /* t0 = A C G T, t1 = A C T T, t2 = A G G - */
    static const uint8_t msa[3][2] = {{1 | 2 << 4, 4 | 8 << 4}, {1 | 2 << 4, 8 | 8 << 4}, {1 | 4 << 4, 4 | 0 << 4}};
    static const uint8_t consensus[4] = {1, 2, 4, 8}; /* first non-gap character per column, as the reference takes it */
    pmfs_tree tree;
    pmfs_ctx* ctx = NULL;
    pmfs_mut* muts = NULL;
    int64_t n = 0, i;
    tree.n_nodes = 5;
    tree.parent = parent;
    tree.child_off = child_off;
    tree.child_idx = child_idx;
    tree.leaf_row = leaf_row;
    if (pmfs_create(&ctx, &tree, 0) != PMFS_OK) {
        fprintf(stderr, "pmfs_create: %s\n", pmfs_last_error());
        return 2;
    }
    if (pmfs_run(ctx, PMFS_FITCH, &msa[0][0], 2, 3, 4, consensus, NULL, &muts, &n) != PMFS_OK) {
        fprintf(stderr, "pmfs_run: %s\n", pmfs_last_error());
        pmfs_destroy(ctx);
        return 3;
    }
    for (i = 0; i < n; i++) /* sorted by (node, pos); type: PMFS_NS / PMFS_NI / PMFS_ND */
        printf("node %d pos %d type %d code %d\n", (int)muts[i].node, (int)muts[i].pos, (int)muts[i].type, (int)muts[i].code);
    pmfs_destroy(ctx);
    return 0;
}
