/* C ABI over the host C++ mirror of the reference's MSA construction path (panman_b200/csrc/host/msa_tree.hpp).
 * C++ callers use pmfs_host::Tree directly; these entry points exist for non-C++ callers and the tests.
 * Paths in comments are relative to /root/reference. */
#ifndef PMFS_HOST_H
#define PMFS_HOST_H

#include <stdint.h>

#include "pmfs.h"

#ifdef __cplusplus
extern "C" {
#endif

/* NucMut (src/panman.hpp:55-61) with mutInfo widened to 32 bits so the C layout has no padding. */
typedef pmfs_host_nucmut_t pmfs_host_nucmut;

const char* pmfs_host_last_error(void);
void pmfs_host_free(void* p);

/* createTreeFromNewickString (src/panman.cpp:300-440) alone; needs no device.  Returns the node count (>0) or a
 * negative status.  parent / n_children: [count] or NULL; names: '\n'-joined identifiers in creation (pre-) order. */
int pmfs_host_parse_newick(const char* newick, int32_t* parent, int32_t* n_children, char* names, int64_t names_cap);

/* Tree::Tree(fin, secondFin, FILE_TYPE::MSA | MSA_OPTIMIZE, reference) (src/panman.hpp:458-459; src/panman.cpp:1264-1639)
 * with the column loop on CUDA device `device`.  low_mem selects MSA_OPTIMIZE (--low-mem-mode, Sankoff).
 * Outputs are malloc'd; release each with pmfs_host_free. */
int pmfs_host_build(const char* msa_path, const char* nwk_path, const char* reference, int low_mem, int device,
                    int32_t* n_nodes, int32_t** parent, char** names, int64_t* n_cols, char** consensus, pmfs_mut** muts,
                    int64_t* n_muts, int64_t** nuc_off, pmfs_host_nucmut** nuc, int64_t* n_nuc, uint32_t** block_words,
                    int64_t* n_block_words);

/* Per-node sort order + greedy <=6-run packing (src/panman.cpp:1435-1456, NucMut range ctor src/panman.hpp:86-128)
 * of records already sorted by (node,pos); needs no device. */
int pmfs_host_pack(const pmfs_mut* muts, int64_t n_muts, int32_t n_nodes, int64_t** nuc_off, pmfs_host_nucmut** nuc,
                   int64_t* n_nuc);

/* .panman output: writePanMAN (src/panmanUtils.cpp:269-297) over TreeGroup::writeToFile (src/panman.cpp:7027-7161) and
 * Tree::getNodesPreorder (:2868-2943) for a tree of this path: Cap'n Proto message per panman.capnp (hand-encoded: one
 * segment, standard stream framing), LZMA level 9 around it when compress != 0 (/usr/bin/xz).
 * pmfs_host_write_panman needs no device: tree in pre-order (parent[], '\n'-joined identifiers, branch lengths or NULL
 * for 0.0), per-node packed NucMut lists as a CSR (as pmfs_host_build / pmfs_pack_last return them), Block consensus
 * words.  pmfs_host_build_panman is the whole `panmanUtils -M msa -N nwk [--reference id] [--low-mem-mode] -o out`
 * (src/panmanUtils.cpp:1487-1543) on CUDA device `device`. */
const char* pmfs_host_writer_last_error(void);
int pmfs_host_write_panman(int32_t n_nodes, const int32_t* parent, const char* names_joined, const float* branch_length,
                           const int64_t* nuc_off, const pmfs_host_nucmut* nuc, const uint32_t* block_words,
                           int64_t n_block_words, const char* out_path, int compress);
int pmfs_host_build_panman(const char* msa_path, const char* nwk_path, const char* reference, int low_mem, int device,
                           const char* out_path, int compress);

#ifdef __cplusplus
}
#endif
#endif /* PMFS_HOST_H */
