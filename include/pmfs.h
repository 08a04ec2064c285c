/* libpmfs -- B200 (sm_100a) per-column small parsimony (Fitch / unit-cost Sankoff) for PanMAN's MSA construction path.
 *
 * This C ABI replaces, in one batched call, the per-column loop the reference runs on the host while a PanMAT is built
 * from an MSA and a Newick tree (paths relative to /root/reference):
 *     Fitch   : src/panman.cpp:1371-1425  -> nucFitch{ForwardPass,BackwardPass,AssignMutations}   src/fitchSankoff.cpp:30,96,131
 *     Sankoff : src/panman.cpp:1558-1603  -> nucSankoff{ForwardPass,BackwardPass,AssignMutations} src/fitchSankoff.cpp:359,487,676
 * Everything before the loop (Newick parse, FASTA read, consensus) and after it (per-node sort, <=6-run NucMut packing,
 * serialisation) stays host C++ (panman_b200/csrc/host/); INTEGRATION.md shows the patch a maintainer would apply.
 *
 * Conventions
 *   - every function returns PMFS_OK (0) or a negative pmfs_status; pmfs_last_error() gives text.  No exceptions, no exit().
 *   - a context is bound to one CUDA device and must be driven by one host thread at a time.
 *   - there is NO CPU fallback: without a usable CUDA device pmfs_create fails with PMFS_ERR_CUDA.
 *   - node ids are the reference's creation order (createTreeFromNewickString, src/panman.cpp:300-440): pre-order,
 *     id 0 is the root ("node_1"), parent[i] < i.
 *   - nucleotide codes are the reference's 4-bit codes (src/panman.cpp:83-118): 0 gap/unknown, A1 C2 M3 G4 R5 S6 V7 T8 W9
 *     Y10 H11 K12 D13 B14 N15; each code is its own unordered state (N is not {A,C,G,T}).
 *   - an MSA is leaf-major, columns contiguous, two columns per byte: column 2j in the LOW nibble of byte j.
 */
#ifndef PMFS_H
#define PMFS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    PMFS_OK = 0,
    PMFS_ERR_ARG = -1,          /* null pointer, bad size, malformed tree */
    PMFS_ERR_CUDA = -2,         /* CUDA runtime error or no device (there is no CPU path) */
    PMFS_ERR_NOMEM = -3,        /* device or host allocation failed */
    PMFS_ERR_ROOT_INVALID = -4, /* Sankoff: a column has no tree leaf present in the MSA and no forced root; the
                                   reference aborts there (assert, src/fitchSankoff.cpp:505) */
    PMFS_ERR_STATE = -5         /* call out of order (e.g. run before an MSA is set) */
} pmfs_status;

enum { PMFS_FITCH = 0, PMFS_SANKOFF = 1 };
enum { PMFS_NS = 0, PMFS_ND = 1, PMFS_NI = 2 }; /* panmanUtils::NucMutationType, src/panman.hpp:26-41 */

/* Tree in the reference's numbering.  Replaces the Node* graph + allNodes map (src/panman.hpp:307-322, 326-340). */
typedef struct {
    int32_t n_nodes;          /* V >= 2; node 0 is the root and must be internal */
    const int32_t* parent;    /* [V], -1 for the root; parent[i] < i */
    const int32_t* child_off; /* [V+1] CSR offsets */
    const int32_t* child_idx; /* children in Newick order (any arity >= 1) */
    const int32_t* leaf_row;  /* [V]: MSA row of a leaf; -1 for internal nodes; -2 for a leaf with no row in the MSA
                                 (the reference gives such a leaf the empty set / all-INF vector,
                                 src/fitchSankoff.cpp:33-36, :363-366) */
} pmfs_tree;

/* One per-branch annotation: the tuple (pos, type, code) the reference pushes into nonGapMutationsMSA[node]
 * (src/panman.cpp:1416-1420, :1597-1601), plus the node it belongs to and the parent's state. */
typedef struct {
    int32_t node;        /* node id */
    int32_t pos;         /* column */
    uint8_t type;        /* PMFS_NS / PMFS_ND / PMFS_NI */
    uint8_t code;        /* getCodeFromNucleotide(mutation char): the node's state, 0 for ND */
    uint8_t parent_code; /* the parent's state (the consensus code for the root); not in the reference tuple */
    uint8_t pad;
} pmfs_mut;

typedef struct pmfs_ctx pmfs_ctx;

/* Timings of the last pmfs_run* call, CUDA events on the context's stream, milliseconds. */
typedef struct {
    float h2d_ms;      /* host->device copies (0 for the device-resident entry points) */
    float forward_ms;  /* bottom-up set pass (all chunk levels) */
    float backward_ms; /* top-down assignment + mutation emission */
    float sort_ms;     /* ordering of the emitted records by (node,pos) */
    float d2h_ms;      /* record download */
    float total_ms;
    int64_t n_launches;     /* kernels of this library launched by the call */
    int64_t forward_bytes;  /* design-level HBM bytes of the forward pass ("bytes per update") */
    int64_t backward_bytes; /* same for the backward pass, excluding emitted records */
    float wall_ms;          /* host wall clock of the device-resident part of the call (launch + sync overheads included) */
    float exchange_ms;      /* pmfs_comm_run: cut points + NCCL exchange of the records by node range */
    float stitch_ms;        /* pmfs_comm_run: owner-side merge of the received slices */
} pmfs_timing;

const char* pmfs_last_error(void);
const char* pmfs_version(void);

/* Builds the per-tree execution plan (chunking + per-chunk walk programs) and uploads it to `device`.
 * Replaces nothing in the reference one-to-one: it is the flat, device-side stand-in for the Node* graph that
 * createTreeFromNewickString (src/panman.cpp:300-440) produces. */
int pmfs_create(pmfs_ctx** ctx, const pmfs_tree* tree, int device);
void pmfs_destroy(pmfs_ctx* ctx);

/* Optional tuning, call before the first run: target nodes per chunk (0 = automatic) and columns per pass
 * (0 = automatic from free HBM).  Results never depend on either. */
int pmfs_set_tuning(pmfs_ctx* ctx, int32_t chunk_nodes, int64_t tile_cols);

/* Column-tile shape of the walk kernels' work items: `walking_warps` x 32 threads, each walking `groups_per_thread`
 * groups of 32 columns.  Built shapes: 4x2 (256 groups per item), 2x2 and 4x1 (128), 1x2 and 2x1 (64), 1x1 (32).
 * 0 = automatic: the shape with the smallest estimated time for the tree and the column range (a tile's cost grows
 * sub-linearly with its width and hardly shrinks when it is partly empty), the narrowest one that still fills the GPU
 * on ladder-like trees (there the column tiles are the only independent work), one group per thread on trees with
 * polytomies.  Results never depend on it.  No counterpart in the reference. */
int pmfs_set_tile(pmfs_ctx* ctx, int32_t walking_warps, int32_t groups_per_thread);

/* Polytomies: a node with MORE than `min_arity` children (default 128; 0 = never) is split -- its children are dealt
 * out, `part_children` at a time (default and maximum 63; 0 = default), to work items of their own that fold them side
 * by side into partial results (Fitch: the AND and the OR of the children's sets, the two things src/fitchSankoff.cpp:39-55
 * needs of them; Sankoff: bit-sliced counts of the children whose set holds each state, :372-403), which the node then
 * combines; the same items walk their children top-down.  Without it a node with thousands of children -- common in
 * outbreak trees -- is one serial stretch per column tile.  The next run re-plans.  Results never depend on it.  No
 * counterpart in the reference. */
int pmfs_set_poly_split(pmfs_ctx* ctx, int32_t min_arity, int32_t part_children);

/* Every wait inside the walk kernels gives up after `seconds` (default 10; 0 = never) and the run fails with
 * PMFS_ERR_CUDA and a message naming the watchdog: a bookkeeping bug must not hang a GPU, but a debugger, a sanitizer
 * or a time-sliced GPU can make legitimate waits long.  A tripped watchdog leaves the CUDA context unusable, like any
 * device-side trap.  No counterpart in the reference. */
int pmfs_set_watchdog(pmfs_ctx* ctx, double seconds);

/* Run all of the context's work on `cuda_stream` (a cudaStream_t of the context's device; NULL restores the context's
 * own stream).  Lets a caller order the passes after its own producers and time them with its own events. */
int pmfs_set_stream(pmfs_ctx* ctx, void* cuda_stream);

/* The reference-facing call: host buffers in, host records out.
 * Replaces the whole column loop src/panman.cpp:1371-1425 (algo = PMFS_FITCH) or :1558-1603 (PMFS_SANKOFF).
 *   msa_nibbles       [n_rows][row_stride_bytes]  leaf-major packed codes (see top of file)
 *   consensus_code    [n_cols]  code compared against the root's state (1<<code(consensusSeq[i]) in :1414-1415,
 *                               code(consensusSeq[i]) in :1590-1596)
 *   root_forced_code  [n_cols] or NULL: Fitch refState (src/fitchSankoff.cpp:45-47, passed at src/panman.cpp:1409-1410
 *                               when --reference is given) / Sankoff defaultValue (:491-492, src/panman.cpp:1573-1594)
 *   out, n_out        records sorted by (node,pos): exactly the order the per-node std::sort of :1437 produces.
 *                     They live in a pinned buffer owned by the context: valid until the next pmfs_run on this context,
 *                     pmfs_result_free or pmfs_destroy.
 * Error behaviour: where the reference exit()s or asserts, this returns a status (PMFS_ERR_ROOT_INVALID, PMFS_ERR_ARG). */
int pmfs_run(pmfs_ctx* ctx, int algo, const uint8_t* msa_nibbles, int64_t row_stride_bytes, int32_t n_rows,
             int64_t n_cols, const uint8_t* consensus_code, const uint8_t* root_forced_code, pmfs_mut** out,
             int64_t* n_out);
void pmfs_result_free(pmfs_ctx* ctx, pmfs_mut* out);

/* Same call on CHARACTERS, as the reference holds them (std::string rows, src/panman.cpp:1271-1315): rows of n_cols
 * alignment characters, consensus and (optional) forced-root characters.  The characters are uploaded and converted to
 * packed codes on the device with the reference's table getCodeFromNucleotide (src/panman.cpp:83-118; case-sensitive:
 * '-', lower case and anything else is code 0), replacing the per-cell conversions of :1398-1406 / :1565-1571. */
int pmfs_run_chars(pmfs_ctx* ctx, int algo, const char* msa_chars, int64_t row_stride_bytes, int32_t n_rows,
                   int64_t n_cols, const char* consensus_chars, const char* root_forced_chars, pmfs_mut** out,
                   int64_t* n_out);

/* The same on rows that are NOT contiguous: rows[r] points to the n_cols characters of sequence r, wherever the caller
 * keeps them -- the reference keeps a std::map<std::string, std::string> (src/panman.cpp:1271-1315), so a caller can pass
 * the map's strings as they are.  The rows go through two pinned staging buffers: every character is touched once on the
 * host, the upload of one chunk of rows overlaps the staging of the next, and the device keeps packed codes only. */
int pmfs_run_char_rows(pmfs_ctx* ctx, int algo, const char* const* rows, int32_t n_rows, int64_t n_cols,
                       const char* consensus_chars, const char* root_forced_chars, pmfs_mut** out, int64_t* n_out);

/* Same call on unpacked STATE CODES, one byte per cell (low 4 bits used): for the callers of the path that do not come
 * from alignment characters.  The reference's block-level passes are this call on three states:
 * blockFitchForwardPassNew / BackwardPassNew / AssignMutationsNew (src/fitchSankoff.cpp:224-308, driven per block at
 * src/panman.cpp:865-904 and src/reroot.cpp:80-82) with PMFS_FITCH, blockSankoffForwardPass / BackwardPass /
 * AssignMutations (:707-818, driven at src/panman.cpp:906-949) with PMFS_SANKOFF, on codes 0 = block absent,
 * 1 = forward strand, 2 = reverse strand, one column per block, consensus_code = all zeros (the root's parent state is
 * "absent"), root_forced_code = the reference sequence's states (defaultState) or NULL.  Records translate with
 * pmfs_block_type / pmfs_block_inverted below.  One difference to mind: the block Sankoff pass gives a leaf without a
 * sequence the state "absent" (src/fitchSankoff.cpp:711-714), so such leaves need an all-zero row rather than
 * leaf_row = -2; block Fitch treats them like the nucleotide pass (empty set).  tests/test_oracle.py pins this against
 * the reference's own block members. */
int pmfs_run_codes(pmfs_ctx* ctx, int algo, const uint8_t* codes, int64_t row_stride_bytes, int32_t n_rows,
                   int64_t n_cols, const uint8_t* consensus_code, const uint8_t* root_forced_code, pmfs_mut** out,
                   int64_t* n_out);
/* BlockMutationType of a record computed on block states (src/panman.hpp:43-52): 1 = BI (insertion), 0 = BD
 * (deletion, or inversion when pmfs_block_inverted is set) -- the pair blockFitchAssignMutationsNew stores. */
static inline int pmfs_block_type(const pmfs_mut* m) { return m->type == 2 /* PMFS_NI */ ? 1 : 0; }
static inline int pmfs_block_inverted(const pmfs_mut* m) { return m->type == 2 ? m->code == 2 : m->type == 0 /* PMFS_NS */; }

/* Device-resident variant (what `value` in bench.py times): every pointer is a DEVICE pointer on the context's
 * device; msa row stride must be a multiple of 16 bytes and the base 16-byte aligned.  Records stay on the device,
 * sorted by (node,pos); *d_out is owned by the context and valid until the next run. */
int pmfs_run_device(pmfs_ctx* ctx, int algo, const uint8_t* d_msa_nibbles, int64_t row_stride_bytes, int32_t n_rows,
                    int64_t n_cols, const uint8_t* d_consensus_code, const uint8_t* d_root_forced_code,
                    pmfs_mut** d_out, int64_t* n_out);
/* ---- column batches --------------------------------------------------------------------------------------------------
 * The same pass over [n_cols] columns in batches of `batch_cols` columns (a multiple of 32; 0 = automatic: as many as
 * the per-node set arrays of one batch fit in free device memory, halved again if one batch would exceed 2^31
 * records).  The reference batches columns the same way when memory is short: --low-mem-mode reads and processes
 * 20 000 columns at a time and appends every batch's tuples to the per-node lists (src/panman.cpp:1503-1514, :1597-1601).
 * After each batch `consumer` (may be NULL) is called with the batch's records ON THE DEVICE, sorted by (node,pos),
 * positions global; the buffer is reused by the next batch, so the consumer must be done with it -- or have queued
 * its copy on the context's stream -- when it returns; a non-zero return value aborts the run.  A node's complete list
 * is the concatenation of its slices from all batches in batch order.  This is how an 8 M-taxon x 30 000-column
 * alignment (2.4 x 10^9 records, 120 GB of nibbles) runs on one GPU.  pmfs_get_timing reports the sums over batches. */
typedef int (*pmfs_batch_fn)(void* user, int64_t col_begin, int64_t n_cols, const pmfs_mut* d_records, int64_t n_records);
int pmfs_run_batched(pmfs_ctx* ctx, int algo, const uint8_t* d_msa_nibbles, int64_t row_stride_bytes, int32_t n_rows,
                     int64_t n_cols, const uint8_t* d_consensus_code, const uint8_t* d_root_forced_code, int64_t batch_cols,
                     pmfs_batch_fn consumer, void* user, int64_t* n_total, int32_t* n_batches);
/* Batch size used by pmfs_run_batched when its batch_cols is 0, and by pmfs_run / pmfs_run_chars / pmfs_run_codes, which
 * stream an alignment that does not fit on the device next to its set arrays through the GPU in column batches
 * (0 = automatic). */
int pmfs_set_batch_cols(pmfs_ctx* ctx, int64_t batch_cols);

/* ---- column shards across GPUs (one context per GPU; no counterpart in the reference, which is single-process) ------
 * pmfs_set_shard puts a context in shard mode: its columns are [pos_offset, pos_offset+n_cols) of a total_cols-wide
 * alignment, and node ids [node_bounds[r], node_bounds[r+1]) are owned by rank r (n_parts <= 16; n_parts = 0 leaves
 * shard mode).  In shard mode pmfs_run_device emits global positions, does NOT order the records and returns
 * *d_out = NULL.  The records (one opaque 64-bit word each) are grouped by owning rank: pmfs_get_parts gives the base
 * device pointer and, per owner r, offsets[r] and counts[r] in records (owner r's block is
 * d_records[offsets[r] .. offsets[r]+counts[r])) for an all-to-all.  *column_ordered is 1 when every block is laid out
 * in column-region order (always, unless ranks x column count is enormous).  The owner concatenates the blocks it
 * received IN SOURCE-RANK ORDER and orders them with pmfs_sort_records (d_records is overwritten; the result is valid
 * until the next run): with column_ordered = 1 on every source a stable sort on the node bits suffices, pass 0 to force
 * a full sort.  The result is the owner's nodes' complete (node,pos)-sorted lists.
 * With n_parts = 0 and total_cols > 0 the context is a column shard that keeps its output: the run orders its own
 * records by (node,pos) as usual, positions are global (pos_offset added), and nothing needs exchanging -- a node's
 * complete list is the concatenation of the shards' slices for that node in column order (pmfs_pack_last gives the
 * per-node offsets).  total_cols = 0 leaves shard mode altogether. */
int pmfs_set_shard(pmfs_ctx* ctx, int64_t pos_offset, int64_t total_cols, int32_t n_parts, const int32_t* node_bounds);
int pmfs_get_parts(pmfs_ctx* ctx, const void** d_records, int64_t* offsets, int64_t* counts, int* column_ordered);
int pmfs_sort_records(pmfs_ctx* ctx, void* d_records, int64_t n, int column_ordered, pmfs_mut** d_out);

/* ---- column shards gathered by node range with NCCL (BASELINE north_star: "Each GPU runs the full tree over its column
 * range, and the per-node mutation lists are gathered with NCCL over NVLink").  No counterpart in the reference, which
 * is a single process on host cores.
 * A pmfs_comm joins one context (= one GPU) to `world` ranks.  Multi-process use (one process per GPU, e.g. under
 * torchrun or MPI): rank 0 calls pmfs_comm_unique_id and ships the 128 bytes to the others by whatever means the job
 * has, then every rank calls pmfs_comm_create (ncclCommInitRank underneath; NCCL is loaded at run time, world = 1 needs
 * none).  Node ids are split evenly: rank r owns nodes [V*r/world, V*(r+1)/world) (pmfs_comm_node_range).
 * pmfs_comm_run is collective: every rank runs the pass over ITS columns [col_offset, col_offset+n_cols) of a
 * total_cols-wide alignment (n_cols = 0 is allowed: the rank only receives), the (node,pos)-ordered 64-bit records are
 * cut at the owners' node boundaries, exchanged in one NCCL group (8 bytes per record) and stitched by the owner: the
 * ranks' column ranges must be disjoint and ascending with the rank, so a node's list is the concatenation of its runs in
 * rank order -- no sort.  On return *d_out (device, owned by the context, valid until the next run) holds the COMPLETE
 * lists of the rank's node range over all ranks' columns, ordered by (node,pos), global positions; pmfs_download and
 * pmfs_pack_last apply to it (the <=6-run packing must come after the gather: runs straddle shard boundaries).  An
 * alignment that does not fit is run as several calls over ascending column windows, each window split over the ranks
 * (pmfs_comm_batch_cols gives the columns per rank and call that fit on every GPU); a node's list is then the
 * concatenation over the calls.  If a rank fails, the others may block in NCCL: treat an error as fatal for the job. */
typedef struct pmfs_comm pmfs_comm;
int pmfs_comm_unique_id(void* id128);
int pmfs_comm_create(pmfs_comm** comm, pmfs_ctx* ctx, const void* id128, int rank, int world);
void pmfs_comm_destroy(pmfs_comm* comm);
int pmfs_comm_node_range(pmfs_comm* comm, int rank, int32_t* node_begin, int32_t* node_end);
int pmfs_comm_batch_cols(pmfs_comm* comm, int algo, int64_t cols_per_rank, int64_t bytes_per_col, int64_t* batch_cols);
/* How the records travel between the ranks of this communicator: 0 = ncclSend / ncclRecv, 1 = device-to-device copies
 * into the owners' inboxes mapped through CUDA IPC (the default where every rank can map every peer; NCCL then only
 * carries cut points and barriers; PMFS_COMM_TRANSPORT=nccl forces 0), 2 = copies inside one process
 * (pmfs_comm_create_local).  -1: null communicator. */
int pmfs_comm_transport(pmfs_comm* comm);
int pmfs_comm_run(pmfs_comm* comm, int algo, const uint8_t* d_msa_nibbles, int64_t row_stride_bytes, int32_t n_rows,
                  int64_t n_cols, int64_t col_offset, int64_t total_cols, const uint8_t* d_consensus_code,
                  const uint8_t* d_root_forced_code, pmfs_mut** d_out, int64_t* n_out);
/* The same ranks as contexts of ONE process (same or different GPUs), exchanging by device-to-device copies at host
 * barriers instead of NCCL: every rank must then be driven by its own host thread. */
int pmfs_comm_create_local(pmfs_comm** comms, pmfs_ctx* const* ctxs, int world);

/* One process, several GPUs -- what a C++ panmanUtils links against (src/panman.cpp:1264, :1457: the MSA-driven Tree
 * constructors): same contract as pmfs_run (host alignment in, ONE (node,pos)-ordered host array out, valid until the
 * next call), with the columns spread over n_gpus devices (devices = NULL: 0 .. n_gpus-1), one host thread and one
 * context per device, ncclCommInitAll underneath.  window_cols > 0 caps the columns in flight over all GPUs
 * (0 = as many as fit). */
typedef struct pmfs_multi pmfs_multi;
int pmfs_multi_create(pmfs_multi** m, const pmfs_tree* tree, int n_gpus, const int* devices);
void pmfs_multi_destroy(pmfs_multi* m);
int pmfs_multi_run(pmfs_multi* m, int algo, const uint8_t* msa_nibbles, int64_t row_stride_bytes, int32_t n_rows,
                   int64_t n_cols, const uint8_t* consensus_code, const uint8_t* root_forced_code, int64_t window_cols,
                   pmfs_mut** out, int64_t* n_out);

/* Copies device records of the last pmfs_run_device / pmfs_sort_records / pmfs_comm_run to a caller buffer of at least n records. */
int pmfs_download(pmfs_ctx* ctx, pmfs_mut* host_out, int64_t n);

/* Fitch pass that allocates nothing.  This entry point has no counterpart in the reference.
 *
 * It is meant for callers that cannot ask the CUDA allocator for memory when they need a result: a latency-bound service
 * loop (cudaMalloc and cudaFree stall the whole device), a process whose HBM already belongs to another allocator, or
 * code running where the allocator must not be called.  It computes exactly what pmfs_run_device(ctx, PMFS_FITCH, ...)
 * computes, on the plan the context already holds, but every intermediate (the set array shared by the two walk
 * kernels, parent states, bit-sliced consensus, unsorted records, sort buffers) comes out of one 32 MiB array with
 * static storage duration in device memory, which every GPU that loads this library keeps until the process exits.
 *   inputs   device pointers as for pmfs_run_device (all columns are processed in one pass, no tiling)
 *   outputs  d_out: caller's device buffer of out_capacity records, filled sorted by (node,pos); *n_out: record count
 * Limits: the call is rejected with PMFS_ERR_ARG and an explanatory pmfs_last_error() when the pieces carved out of
 * the array, each rounded up to 256 bytes -- with G = ceil(n_cols/32): the set array n_set_rows*G*64 (n_set_rows = n_internal, plus 2 rows per part of a split polytomy,
 * pmfs_get_plan_info), the chunk parent
 * states n_chunks*G*20, two bit-sliced code arrays G*16 each, two record buffers out_capacity*8 each, the radix sort's
 * workspace (about 1/16 of a record buffer) and the walk kernels' counters (64 + 2*n_chunks*tiles)*4 -- exceed 32 MiB,
 * and when more than out_capacity records are produced (*n_out then holds the number needed).  A context in
 * column-shard mode is refused with PMFS_ERR_STATE, and so is one whose tree has split polytomies and whose plan was
 * last built for Sankoff (re-planning would allocate: run pmfs_run_device(PMFS_FITCH) once first).  The call returns after the context's stream has been
 * synchronised; pmfs_download and pmfs_pack_last do not apply to its result (it is in the caller's buffer).
 * The first call on a device looks up the array's address on that device; make that first call before any CUDA stream
 * capture begins.  All calls share the array: it keeps the previous call's contents (the call clears the two words
 * that must start at zero), and two calls whose GPU work overlaps -- different contexts or streams, same device --
 * race on it.  No cudaMalloc, cudaFree or other allocating routine is called. */
int pmfs_run_fitch_static(pmfs_ctx* ctx, const uint8_t* d_msa_nibbles, int64_t row_stride_bytes, int32_t n_rows,
                          int64_t n_cols, const uint8_t* d_consensus_code, const uint8_t* d_root_forced_code,
                          pmfs_mut* d_out, int64_t out_capacity, int64_t* n_out);

/* Greedy <=6-run packing of the last result (pmfs_run / pmfs_run_device / pmfs_sort_records) ON THE DEVICE, followed by
 * the download of the packed lists.  Replaces the per-node loop of src/panman.cpp:1439-1455 with the NucMut range
 * constructor src/panman.hpp:86-128: only 24 bytes per up-to-6 records cross PCIe instead of 12 per
 * record.  *out: n_out packed records, all nodes concatenated; node v's list is out[node_off[v] .. node_off[v+1]).
 * Both arrays live in pinned buffers owned by the context, valid until the next call. */
typedef struct {
    int32_t nucPosition, nucGapPosition, primaryBlockId, secondaryBlockId;
    uint32_t mutInfo; /* uint8_t in the reference's NucMut (src/panman.hpp:55-61); widened for a padding-free layout */
    uint32_t nucs;
} pmfs_host_nucmut_t;
int pmfs_pack_last(pmfs_ctx* ctx, const pmfs_host_nucmut_t** out, int64_t* n_out, const int64_t** node_off);

int pmfs_get_timing(pmfs_ctx* ctx, pmfs_timing* t);

/* Plan introspection (tests, DESIGN.md numbers). */
typedef struct {
    int32_t n_nodes, n_internal, n_leaves;
    int32_t n_chunks, n_fwd_levels, n_bwd_levels;
    int32_t fwd_stack_depth, bwd_stack_depth;
    int32_t max_arity;
    int64_t program_words;
    int64_t edges_leaf;     /* parent->leaf edges */
    int64_t edges_internal; /* parent->internal edges */
    int32_t n_parts;        /* part chunks of split polytomies (counted in n_chunks) */
    int32_t n_set_rows;     /* rows of the set arrays: n_internal + the parts' partial results */
} pmfs_plan_info;
int pmfs_get_plan_info(pmfs_ctx* ctx, pmfs_plan_info* info);

/* ---- synthetic alignments, generated on the device (bench and large-size tests) --------------------------
 * Simulated evolution down the context's tree; a pure function of (seed, node, column), so any column range can be
 * generated independently (column shards on different GPUs see the same global alignment). */
typedef struct {
    uint64_t seed;
    float p_sub, p_del, p_ins; /* per branch, per column */
    float p_n_block;           /* per leaf, per 32-column block: whole block becomes N */
    float p_amb;               /* per leaf cell: one of the 10 two/three-base IUPAC codes */
    float accessory_frac;      /* fraction of columns whose root state is gap */
} pmfs_synth_spec;

/* Fills d_msa_nibbles ([n_leaf_rows][row_stride_bytes], device) with columns [col_begin, col_begin+n_cols) of the
 * synthetic alignment and d_consensus_code ([n_cols], device) with the first non-gap code in row order
 * (the rule of src/panman.cpp:1326-1340 with rows in id order).  Every column has at least one non-gap cell. */
int pmfs_synth_msa(pmfs_ctx* ctx, const pmfs_synth_spec* spec, int64_t col_begin, int64_t n_cols,
                   uint8_t* d_msa_nibbles, int64_t row_stride_bytes, uint8_t* d_consensus_code);

#ifdef __cplusplus
}
#endif
#endif /* PMFS_H */
