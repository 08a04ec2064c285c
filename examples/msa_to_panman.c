/* `panmanUtils -M <msa.fa> -N <tree.nwk> [--reference <id>] [--low-mem-mode] -o <out>` for the path this library
 * covers (src/panmanUtils.cpp:1487-1543 in the reference): FASTA + Newick in, <out>.panman out, the Fitch / Sankoff pass
 * on the GPU.
 *   gcc -std=c99 -Iinclude examples/msa_to_panman.c -Lpanman_b200 -lpmfs -Wl,-rpath,$PWD/panman_b200 -o msa_to_panman
 *   ./msa_to_panman test/sars_20.fa test/sars_20.nwk out [--reference ID] [--low-mem-mode] [--device N] [--no-xz]
 * writes out.panman (xz-compressed Cap'n Proto, as writePanMAN does). */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "pmfs_host.h"

int main(int argc, char** argv) {
    const char* reference = "";
    int low_mem = 0, device = 0, compress = 1, i, rc;
    char out_path[4096];
    if (argc < 4) {
        fprintf(stderr, "usage: %s <msa.fa> <tree.nwk> <out> [--reference ID] [--low-mem-mode] [--device N] [--no-xz]\n", argv[0]);
        return 64;
    }
    for (i = 4; i < argc; i++) {
        if (!strcmp(argv[i], "--reference") && i + 1 < argc) reference = argv[++i];
        else if (!strcmp(argv[i], "--low-mem-mode")) low_mem = 1;
        else if (!strcmp(argv[i], "--device") && i + 1 < argc) device = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--no-xz")) compress = 0;
        else {
            fprintf(stderr, "unknown option %s\n", argv[i]);
            return 64;
        }
    }
    snprintf(out_path, sizeof out_path, "%s.panman", argv[3]);
    rc = pmfs_host_build_panman(argv[1], argv[2], reference, low_mem, device, out_path, compress);
    if (rc != 0) {
        const char* e = pmfs_host_last_error();
        fprintf(stderr, "failed (%d): %s%s%s\n", rc, e ? e : "", e && *e ? "; " : "", pmfs_host_writer_last_error());
        return 1;
    }
    printf("wrote %s\n", out_path);
    return 0;
}
