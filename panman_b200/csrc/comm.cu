// libpmfs across the GPUs of one box: column shards, per-node mutation lists gathered by node range with NCCL over
// NVLink (BASELINE.json north_star: "Each GPU runs the full tree over its column range, and the per-node
// mutation lists are gathered with NCCL").  No counterpart in the reference, which is one process on host cores.
//
// Data path of one call (pmfs_comm_run), rank r of N:
//   1. the ordinary pass over this rank's columns (pmfs.cu run_device_impl) with the ordered result left as 64-bit records
//      (node, global column, payload) plus, from the same prefix sums that ordered them, the first record of every node
//      (noff) and of every rank's node range (cut): the records a rank owns are one contiguous slice of every source;
//   2. ncclAllGather of the N+1 cut points (the only host synchronisation of the exchange: NCCL wants counts on the host);
//   3. ONE NCCL group: the slices (8 bytes per record) and the matching noff ranges (4 bytes per owned node) to their owners;
//   4. the owner stitches: every source's slice is (node,pos)-ordered and the sources' columns are disjoint and
//      rank-ordered, so node v's list is the concatenation of its runs in source order -- per-node totals, one prefix sum,
//      one copy kernel that also turns the records into pmfs_mut.  No sort anywhere.
// Step 3 has two transports.  NCCL send/recv (PMFS_COMM_TRANSPORT=nccl) moved 250 GB/s between two B200s and less with
// more peers -- a third of a step at 8 GPUs.  The default for ranks in different processes of one box is therefore
// PEER COPIES: every rank's inbox is exported once with CUDA IPC, and a source copies each slice straight into its
// owner's inbox at the place the gathered cut points assign to it (copy engines over NVLink, no SMs, no staging); a tiny
// NCCL all-reduce after the copies is the stream-ordered barrier that tells every owner its inbox is complete.  NCCL
// stays the control plane (cut points, barriers, handle exchange).  Inboxes grow collectively (every rank derives every
// rank's need from the same cut points), so an exported buffer is never freed while a peer has it mapped.
// NCCL is loaded at run time (dlopen), so single-GPU users of the library do not need it installed.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <cub/device/device_scan.cuh>

#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>

#include "ctx.hpp"

namespace {

struct NcclApi {
    void* lib = nullptr;
    std::string err;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommInitAll) CommInitAll = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclSend) Send = nullptr;
    decltype(&ncclRecv) Recv = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
};

NcclApi& nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        // a process that already uses NCCL (torch ships its own copy under this soname) gets that one
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            api.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (api.lib) break;
        }
        if (!api.lib) {
            api.err = std::string("cannot load NCCL (libnccl.so.2): ") + dlerror();
            return;
        }
        auto sym = [&](const char* n) -> void* {
            void* p = dlsym(api.lib, n);
            if (!p && api.err.empty()) api.err = std::string("NCCL symbol missing: ") + n;
            return p;
        };
        api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
        api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
        api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
        api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
        api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
        api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
        api.Send = (decltype(api.Send))sym("ncclSend");
        api.Recv = (decltype(api.Recv))sym("ncclRecv");
        api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
        api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
    });
    return api;
}

#define PMFS_NCCL(call)                                                                       \
    do {                                                                                      \
        ncclResult_t r_ = (call);                                                             \
        if (r_ != ncclSuccess) {                                                              \
            g_error = std::string(#call) + ": " + nccl_api().GetErrorString(r_);              \
            return PMFS_ERR_CUDA;                                                             \
        }                                                                                     \
    } while (0)

constexpr int kMaxRanks = 16;

struct StitchParams {
    const unsigned long long* inbox;  // the sources' slices, concatenated in source-rank order
    const uint32_t* noff_in;          // [world][n_own + 1] first record of every owned node in every source's ordering
    unsigned long long roff[kMaxRanks];  // where source s's slice starts in the inbox
    int32_t world, n_own, node_base;
    int pos_bits;
};

// records of owned node v over all sources
__global__ void stitch_count_kernel(const StitchParams p, uint32_t* __restrict__ total) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= p.n_own) return;
    uint32_t t = 0;
    for (int s = 0; s < p.world; s++) {
        const uint32_t* o = p.noff_in + (int64_t)s * (p.n_own + 1) + v;
        t += o[1] - o[0];
    }
    total[v] = t;
}

// one warp per owned node: its runs, in source order, to out[start[v] ...], as pmfs_mut
__global__ void stitch_copy_kernel(const StitchParams p, const uint32_t* __restrict__ start, pmfs_mut* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const unsigned long long pos_mask = (1ull << p.pos_bits) - 1;
    uint32_t* const out32 = (uint32_t*)out;
    for (int64_t v = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; v < p.n_own; v += n_warps) {
        size_t at = start[v];
        for (int s = 0; s < p.world; s++) {
            const uint32_t* o = p.noff_in + (int64_t)s * (p.n_own + 1);
            const uint32_t b = o[v] - o[0], cnt = o[v + 1] - o[v];
            const unsigned long long* src = p.inbox + p.roff[s] + b;
            for (uint32_t i = lane; i < cnt; i += 32) {
                const unsigned long long r = src[i], key = r >> pmfs::kPayloadBits;
                uint32_t* q = out32 + (at + i) * 3;
                q[0] = (uint32_t)(key >> p.pos_bits);
                q[1] = (uint32_t)(key & pos_mask);
                q[2] = (uint32_t)(r & 3) | ((uint32_t)((r >> 2) & 15) << 8) | ((uint32_t)((r >> 6) & 15) << 16);
            }
            at += cnt;
        }
    }
}

}  // namespace

namespace {

struct Barrier {  // reusable barrier for host threads
    std::mutex mu;
    std::condition_variable cv;
    int count = 0, n = 0, gen = 0;
    explicit Barrier(int n_) : n(n_) {}
    void wait() {
        std::unique_lock<std::mutex> lk(mu);
        const int g = gen;
        if (++count == n) {
            count = 0;
            gen++;
            cv.notify_all();
        } else {
            cv.wait(lk, [&] { return gen != g; });
        }
    }
};

// In-process transport: the ranks are contexts of one process, each driven by its own host thread; they meet at a host
// barrier and every owner PULLS its slices from the sources' buffers with device-to-device copies on its own stream
// (copy engines over NVLink between GPUs with peer access; plain copies when the contexts share a GPU, which is how
// the exchange and stitch logic is tested on a single GPU).
struct LocalGroup {
    int world;
    Barrier bar;
    std::vector<pmfs_comm*> members;
    unsigned long long cuts[kMaxRanks][kMaxRanks + 1] = {};
    const unsigned long long* skeys[kMaxRanks] = {};
    const uint32_t* noff[kMaxRanks] = {};
    long long red[kMaxRanks] = {};
    int refs;
    explicit LocalGroup(int w) : world(w), bar(w), members((size_t)w, nullptr), refs(w) {}
};

}  // namespace

struct pmfs_comm {
    pmfs_ctx* ctx = nullptr;
    ncclComm_t comm = nullptr;
    LocalGroup* group = nullptr;  // in-process transport instead of NCCL
    int rank = 0, world = 1;
    PartBounds parts = {};
    DevBuf<unsigned long long> d_allcuts, d_inbox;
    unsigned long long* h_allcuts = nullptr;  // pinned [world][world + 1]
    DevBuf<uint32_t> d_noff_in, d_total, d_start;
    DevBuf<uint8_t> d_scan_tmp;
    DevBuf<long long> d_red;
    long long* h_red = nullptr;  // pinned
    cudaEvent_t ev[3] = {};
    float exchange_ms = 0, stitch_ms = 0;
    // peer-copy transport
    bool p2p = false;
    size_t inbox_cap_all[kMaxRanks] = {};        // every rank's inbox capacity in records (the same on all ranks)
    unsigned long long* peer_inbox[kMaxRanks] = {};  // IPC mappings of the peers' inboxes / node-offset tables
    uint32_t* peer_noff_in[kMaxRanks] = {};
    DevBuf<unsigned long long> d_hand;           // handle exchange: [world][8] words
    unsigned long long* h_hand = nullptr;        // pinned
};

static int p2p_setup(pmfs_comm* m);

static int comm_finish_create(pmfs_comm* m) {
    pmfs_ctx* c = m->ctx;
    const int32_t V = c->plan.n_nodes;
    m->parts.n_parts = m->world;
    for (int r = 0; r <= m->world; r++) m->parts.bound[r] = (int32_t)(((int64_t)V * r) / m->world);
    PMFS_CUDA(cudaSetDevice(c->device));
    PMFS_CUDA(cudaMallocHost(&m->h_allcuts, (size_t)kMaxRanks * (kMaxRanks + 1) * sizeof(unsigned long long)));
    PMFS_CUDA(cudaMallocHost(&m->h_red, 8 * sizeof(long long)));
    for (auto& e : m->ev) PMFS_CUDA(cudaEventCreate(&e));
    int rc;
    if ((rc = m->d_allcuts.ensure((size_t)kMaxRanks * (kMaxRanks + 1)))) return rc;
    if ((rc = m->d_red.ensure(8))) return rc;
    return PMFS_OK;
}

// ---- peer-copy transport: CUDA IPC mappings of the peers' buffers, exchanged through NCCL -----------------------------
// All ranks gather everybody's handle for one buffer (`mine`, or nothing for ranks not in `mask`) and map the ones in mask.
static int exchange_handles(pmfs_comm* m, void* mine, uint32_t mask, void** mapped) {
    pmfs_ctx* c = m->ctx;
    NcclApi& A = nccl_api();
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    const int W = m->world, me = m->rank;
    int rc;
    if ((rc = m->d_hand.ensure((size_t)kMaxRanks * 8 + 8))) return rc;
    if (!m->h_hand) PMFS_CUDA(cudaMallocHost(&m->h_hand, ((size_t)kMaxRanks * 8 + 8) * sizeof(unsigned long long)));
    cudaIpcMemHandle_t h = {};
    bool exported = true;  // a rank that cannot export still takes part in the gather (an all-zero handle maps nowhere)
    if (mine && ((mask >> me) & 1u) && cudaIpcGetMemHandle(&h, mine) != cudaSuccess) {
        g_error = "pmfs_comm: cudaIpcGetMemHandle failed -- set PMFS_COMM_TRANSPORT=nccl";
        cudaGetLastError();
        memset(&h, 0, sizeof(h));
        exported = false;
    }
    memcpy(m->h_hand + (size_t)kMaxRanks * 8, &h, 64);
    PMFS_CUDA(cudaMemcpyAsync(m->d_hand.ptr + (size_t)kMaxRanks * 8, m->h_hand + (size_t)kMaxRanks * 8, 64, cudaMemcpyHostToDevice, c->stream));
    PMFS_NCCL(A.AllGather(m->d_hand.ptr + (size_t)kMaxRanks * 8, m->d_hand.ptr, 8, ncclUint64, m->comm, c->stream));
    PMFS_CUDA(cudaMemcpyAsync(m->h_hand, m->d_hand.ptr, (size_t)W * 64, cudaMemcpyDeviceToHost, c->stream));
    PMFS_CUDA(cudaStreamSynchronize(c->stream));
    for (int r = 0; r < W; r++) {
        if (r == me || !((mask >> r) & 1u)) continue;
        cudaIpcMemHandle_t hr;
        memcpy(&hr, m->h_hand + (size_t)r * 8, 64);
        void* ptr = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&ptr, hr, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            g_error = std::string("pmfs_comm: cudaIpcOpenMemHandle (rank ") + std::to_string(r) + "): " + cudaGetErrorString(e) +
                      " -- set PMFS_COMM_TRANSPORT=nccl where the GPUs cannot map each other's memory";
            cudaGetLastError();
            return PMFS_ERR_CUDA;
        }
        mapped[r] = ptr;
    }
    return exported ? PMFS_OK : PMFS_ERR_CUDA;
}

// stream-ordered barrier over the ranks (a one-word all-reduce); host_sync: also wait for it on the host
static int comm_barrier(pmfs_comm* m, bool host_sync) {
    pmfs_ctx* c = m->ctx;
    PMFS_NCCL(nccl_api().AllReduce(m->d_red.ptr, m->d_red.ptr, 1, ncclInt64, ncclSum, m->comm, c->stream));
    if (host_sync) PMFS_CUDA(cudaStreamSynchronize(c->stream));
    return PMFS_OK;
}

static int p2p_setup(pmfs_comm* m) {
    // the node-offset tables never change size: W x (owned nodes + 1) words, exported once
    pmfs_ctx* c = m->ctx;
    const int W = m->world, me = m->rank;
    const int32_t n_own = m->parts.bound[me + 1] - m->parts.bound[me];
    int rc;
    if ((rc = m->d_noff_in.ensure((size_t)W * ((size_t)n_own + 1)))) return rc;
    const int mapped = exchange_handles(m, m->d_noff_in.ptr, (1u << W) - 1u, (void**)m->peer_noff_in);
    // Either every rank has mapped every peer, or nobody uses the mappings: agree on it (the ranks must not part ways).
    m->h_red[0] = mapped == PMFS_OK ? 0 : 1;
    PMFS_CUDA(cudaMemcpyAsync(m->d_red.ptr, m->h_red, sizeof(long long), cudaMemcpyHostToDevice, c->stream));
    PMFS_NCCL(nccl_api().AllReduce(m->d_red.ptr, m->d_red.ptr, 1, ncclInt64, ncclSum, m->comm, c->stream));
    PMFS_CUDA(cudaMemcpyAsync(m->h_red, m->d_red.ptr, sizeof(long long), cudaMemcpyDeviceToHost, c->stream));
    PMFS_CUDA(cudaStreamSynchronize(c->stream));
    if (m->h_red[0] != 0) {  // some rank cannot map a peer (no peer access, separate IPC namespaces): ncclSend / ncclRecv
        for (int r = 0; r < W; r++)
            if (r != me && m->peer_noff_in[r]) {
                cudaIpcCloseMemHandle(m->peer_noff_in[r]);
                m->peer_noff_in[r] = nullptr;
            }
        cudaGetLastError();
        m->p2p = false;
    }
    return PMFS_OK;
}

// Every rank's inbox must hold what the gathered cut points say it will receive.  Inboxes that have to grow do so
// collectively: everybody unmaps them, a barrier, the owners reallocate and re-export, everybody maps the new ones.
static int p2p_grow_inboxes(pmfs_comm* m, const unsigned long long* all) {
    const int W = m->world, me = m->rank;
    uint32_t grow = 0;
    size_t newcap[kMaxRanks] = {};
    for (int p = 0; p < W; p++) {
        unsigned long long need = 0;
        for (int s = 0; s < W; s++) need += all[s * (W + 1) + p + 1] - all[s * (W + 1) + p];
        if (need > m->inbox_cap_all[p]) {
            grow |= 1u << p;
            newcap[p] = (size_t)need + (size_t)need / 4 + 1024;
        }
    }
    if (!grow) return PMFS_OK;
    for (int p = 0; p < W; p++)
        if (((grow >> p) & 1u) && p != me && m->peer_inbox[p]) {
            PMFS_CUDA(cudaIpcCloseMemHandle(m->peer_inbox[p]));
            m->peer_inbox[p] = nullptr;
        }
    int rc = comm_barrier(m, true);  // nobody has the old buffers mapped any more
    if (rc) return rc;
    if ((grow >> me) & 1u) {
        m->d_inbox.release();
        if ((rc = m->d_inbox.ensure(newcap[me]))) return rc;
    }
    if ((rc = exchange_handles(m, m->d_inbox.ptr, grow, (void**)m->peer_inbox))) return rc;
    for (int p = 0; p < W; p++)
        if ((grow >> p) & 1u) m->inbox_cap_all[p] = newcap[p];
    return PMFS_OK;
}

extern "C" {

int pmfs_comm_unique_id(void* id128) {
    NcclApi& A = nccl_api();
    if (!A.err.empty() || !id128) {
        g_error = !id128 ? "pmfs_comm_unique_id: null argument" : "pmfs_comm_unique_id: " + A.err;
        return PMFS_ERR_CUDA;
    }
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    PMFS_NCCL(A.GetUniqueId(&id));
    memcpy(id128, &id, sizeof(id));
    return PMFS_OK;
}

int pmfs_comm_create(pmfs_comm** out, pmfs_ctx* ctx, const void* id128, int rank, int world) {
    if (!out || !ctx || (!id128 && world > 1) || world < 1 || world > kMaxRanks || rank < 0 || rank >= world) {
        g_error = "pmfs_comm_create: bad argument (at most 16 ranks)";
        return PMFS_ERR_ARG;
    }
    *out = nullptr;
    PMFS_CUDA(cudaSetDevice(ctx->device));
    pmfs_comm* m = new pmfs_comm();
    m->ctx = ctx;
    m->rank = rank;
    m->world = world;
    if (world > 1) {  // (a single rank gathers nothing: NCCL need not even be installed)
        NcclApi& A = nccl_api();
        if (!A.err.empty()) {
            g_error = "pmfs_comm_create: " + A.err;
            delete m;
            return PMFS_ERR_CUDA;
        }
        ncclUniqueId id;
        memcpy(&id, id128, sizeof(id));
        ncclResult_t r = A.CommInitRank(&m->comm, world, id, rank);
        if (r != ncclSuccess) {
            g_error = std::string("ncclCommInitRank: ") + A.GetErrorString(r);
            delete m;
            return PMFS_ERR_CUDA;
        }
    }
    int rc = comm_finish_create(m);
    if (!rc && world > 1) {
        const char* tr = getenv("PMFS_COMM_TRANSPORT");
        m->p2p = !(tr && std::string(tr) == "nccl");
        if (m->p2p) rc = p2p_setup(m);
    }
    if (rc) {
        pmfs_comm_destroy(m);
        return rc;
    }
    *out = m;
    return PMFS_OK;
}

void pmfs_comm_destroy(pmfs_comm* m) {
    if (!m) return;
    cudaSetDevice(m->ctx->device);
    cudaStreamSynchronize(m->ctx->stream);
    for (int r = 0; r < kMaxRanks; r++) {
        if (m->peer_inbox[r]) cudaIpcCloseMemHandle(m->peer_inbox[r]);
        if (m->peer_noff_in[r]) cudaIpcCloseMemHandle(m->peer_noff_in[r]);
    }
    if (m->h_hand) cudaFreeHost(m->h_hand);
    m->d_hand.release();
    if (m->comm) nccl_api().CommDestroy(m->comm);
    if (m->group && --m->group->refs == 0) delete m->group;  // (destroy the members from one thread)
    m->d_allcuts.release(); m->d_inbox.release(); m->d_noff_in.release(); m->d_total.release(); m->d_start.release();
    m->d_scan_tmp.release(); m->d_red.release();
    if (m->h_allcuts) cudaFreeHost(m->h_allcuts);
    if (m->h_red) cudaFreeHost(m->h_red);
    for (auto& e : m->ev)
        if (e) cudaEventDestroy(e);
    delete m;
}

// `world` communicators over contexts of THIS process (same or different GPUs), exchanging through device-to-device
// copies instead of NCCL.  Every rank must be driven by its own host thread: pmfs_comm_run blocks at host barriers.
int pmfs_comm_create_local(pmfs_comm** comms, pmfs_ctx* const* ctxs, int world) {
    if (!comms || !ctxs || world < 1 || world > kMaxRanks) {
        g_error = "pmfs_comm_create_local: bad argument (at most 16 ranks)";
        return PMFS_ERR_ARG;
    }
    LocalGroup* g = new LocalGroup(world);
    for (int r = 0; r < world; r++) {
        pmfs_comm* m = new pmfs_comm();
        m->ctx = ctxs[r];
        m->rank = r;
        m->world = world;
        m->group = g;
        g->members[r] = m;
        comms[r] = m;
        int rc = comm_finish_create(m);
        if (rc) return rc;
    }
    for (int a = 0; a < world; a++)  // peer access where the contexts sit on different GPUs (ignored if unavailable)
        for (int b = 0; b < world; b++)
            if (ctxs[a]->device != ctxs[b]->device) {
                int can = 0;
                cudaDeviceCanAccessPeer(&can, ctxs[a]->device, ctxs[b]->device);
                if (can) {
                    cudaSetDevice(ctxs[a]->device);
                    if (cudaDeviceEnablePeerAccess(ctxs[b]->device, 0) != cudaSuccess) cudaGetLastError();
                }
            }
    return PMFS_OK;
}

int pmfs_comm_node_range(pmfs_comm* m, int rank, int32_t* node_begin, int32_t* node_end) {
    if (!m || rank < 0 || rank >= m->world || !node_begin || !node_end) {
        g_error = "pmfs_comm_node_range: bad argument";
        return PMFS_ERR_ARG;
    }
    *node_begin = m->parts.bound[rank];
    *node_end = m->parts.bound[rank + 1];
    return PMFS_OK;
}

// Columns per rank and call such that every rank's set arrays fit in what its GPU has free: the minimum over ranks
// (one small all-reduce), a multiple of 32 column groups where possible.  `bytes_per_col`: device memory the caller will
// still allocate per column of a call (e.g. n_rows / 2 for a staged alignment window).
int pmfs_comm_transport(pmfs_comm* m) {
    if (!m) return -1;
    return m->group ? 2 : (m->p2p ? 1 : 0);
}

int pmfs_comm_batch_cols(pmfs_comm* m, int algo, int64_t cols_per_rank, int64_t bytes_per_col, int64_t* batch_cols) {
    if (!m || cols_per_rank <= 0 || !batch_cols || bytes_per_col < 0) {
        g_error = "pmfs_comm_batch_cols: bad argument";
        return PMFS_ERR_ARG;
    }
    pmfs_ctx* c = m->ctx;
    PMFS_CUDA(cudaSetDevice(c->device));
    const int64_t G_all = (cols_per_rank + 31) / 32;
    int64_t g = 0;
    int rc = mem_limited_groups(c, algo, G_all, (size_t)bytes_per_col * 32, &g);
    if (rc) return rc;
    const int64_t max_regions = std::max<int64_t>(1, ((int64_t)1 << 30) / std::max(1, c->plan.n_nodes));
    g = std::min(g, max_regions * 32);
    if (m->group) {
        m->group->red[m->rank] = g;
        m->group->bar.wait();
        for (int r = 0; r < m->world; r++) g = std::min<int64_t>(g, m->group->red[r]);
        m->group->bar.wait();
    } else if (m->world > 1) {
        m->h_red[0] = g;
        PMFS_CUDA(cudaMemcpyAsync(m->d_red.ptr, m->h_red, sizeof(long long), cudaMemcpyHostToDevice, c->stream));
        PMFS_NCCL(nccl_api().AllReduce(m->d_red.ptr, m->d_red.ptr, 1, ncclInt64, ncclMin, m->comm, c->stream));
        PMFS_CUDA(cudaMemcpyAsync(m->h_red, m->d_red.ptr, sizeof(long long), cudaMemcpyDeviceToHost, c->stream));
        PMFS_CUDA(cudaStreamSynchronize(c->stream));
        g = m->h_red[0];
    }
    g = balanced_batch_groups(G_all, g);
    *batch_cols = std::min(cols_per_rank, g * 32);
    return PMFS_OK;
}

int pmfs_comm_run(pmfs_comm* m, int algo, const uint8_t* d_msa, int64_t row_stride, int32_t n_rows, int64_t n_cols,
                  int64_t col_offset, int64_t total_cols, const uint8_t* d_cons, const uint8_t* d_forced, pmfs_mut** d_out,
                  int64_t* n_out) {
    if (!m || !n_out || n_cols < 0 || col_offset < 0 || total_cols <= 0 || col_offset + n_cols > total_cols) {
        g_error = "pmfs_comm_run: bad argument";
        return PMFS_ERR_ARG;
    }
    pmfs_ctx* c = m->ctx;
    NcclApi& A = nccl_api();
    PMFS_CUDA(cudaSetDevice(c->device));
    const int W = m->world, me = m->rank;
    // 1. this rank's columns; ordered 64-bit records with global positions, node offsets, owner cuts
    const int64_t keep_off = c->shard_pos_offset, keep_total = c->shard_total_cols;
    const PartBounds keep_parts = c->parts;
    c->parts = PartBounds();
    c->shard_pos_offset = col_offset;
    c->shard_total_cols = total_cols;
    c->keys_out = true;
    c->xparts = m->parts;
    int64_t n_local = 0;
    c->timing.h2d_ms = 0;
    int rc = PMFS_OK;
    if (n_cols > 0) rc = run_device_impl(c, algo, d_msa, row_stride, n_rows, n_cols, d_cons, d_forced, nullptr, &n_local);
    else c->timing = pmfs_timing();  // (a rank without columns in this call still takes part in the exchange)
    c->keys_out = false;
    c->shard_pos_offset = keep_off;
    c->shard_total_cols = keep_total;
    c->parts = keep_parts;
    if (rc) return rc;
    const int32_t V = c->plan.n_nodes;
    if (n_cols == 0) {  // (no pass ran: every node starts at 0)
        if ((rc = c->d_noff.ensure((size_t)V + 1))) return rc;
        if ((rc = c->d_cuts.ensure(32))) return rc;
        if (!c->h_cuts) PMFS_CUDA(cudaMallocHost(&c->h_cuts, 32 * sizeof(unsigned long long)));
        PMFS_CUDA(cudaMemsetAsync(c->d_noff.ptr, 0, ((size_t)V + 1) * sizeof(uint32_t), c->stream));
        PMFS_CUDA(cudaMemsetAsync(c->d_cuts.ptr, 0, 32 * sizeof(unsigned long long), c->stream));
        memset(c->h_cuts, 0, 32 * sizeof(unsigned long long));
        PMFS_CUDA(cudaStreamSynchronize(c->stream));
    }
    PMFS_CUDA(cudaEventRecord(m->ev[0], c->stream));
    // 2. everybody's cut points
    unsigned long long* all = m->h_allcuts;
    LocalGroup* G = m->group;
    if (G) {  // (the pass has synchronised this rank's stream: its records, offsets and h_cuts are final)
        memcpy(G->cuts[me], c->h_cuts, (size_t)(W + 1) * sizeof(unsigned long long));
        G->skeys[me] = c->skeys_view;
        G->noff[me] = c->d_noff.ptr;
        G->bar.wait();
        for (int s = 0; s < W; s++) memcpy(all + (size_t)s * (W + 1), G->cuts[s], (size_t)(W + 1) * sizeof(unsigned long long));
    } else if (W > 1) {
        PMFS_NCCL(A.AllGather(c->d_cuts.ptr, m->d_allcuts.ptr, (size_t)(W + 1), ncclUint64, m->comm, c->stream));
        PMFS_CUDA(cudaMemcpyAsync(all, m->d_allcuts.ptr, (size_t)W * (W + 1) * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
        PMFS_CUDA(cudaStreamSynchronize(c->stream));
    } else {
        memcpy(all, c->h_cuts, 2 * sizeof(unsigned long long));
    }
    StitchParams sp = {};
    unsigned long long total = 0;
    for (int s = 0; s < W; s++) {
        sp.roff[s] = total;
        total += all[s * (W + 1) + me + 1] - all[s * (W + 1) + me];
    }
    if (total > 0x7fffffffull) {
        g_error = "pmfs_comm_run: more than 2^31 records for one rank's node range in one call: run fewer columns per call";
        return PMFS_ERR_ARG;
    }
    const int32_t n_own = m->parts.bound[me + 1] - m->parts.bound[me];
    const bool p2p = m->p2p && !G && W > 1;
    if (p2p) {
        if ((rc = p2p_grow_inboxes(m, all))) return rc;
    } else {
        if ((rc = m->d_inbox.ensure((size_t)total + (size_t)total / 8 + 16))) return rc;
    }
    if ((rc = m->d_noff_in.ensure((size_t)W * ((size_t)n_own + 1)))) return rc;
    if ((rc = m->d_total.ensure((size_t)n_own + 1))) return rc;
    if ((rc = m->d_start.ensure((size_t)n_own + 1))) return rc;
    if ((rc = c->d_out.ensure((size_t)total + (size_t)total / 8 + 16))) return rc;
    // 3. slices and node offsets to their owners: one NCCL group; this rank's own slice by a device copy
    const unsigned long long* mine = c->skeys_view;
    const unsigned long long* my_cuts = all + (size_t)me * (W + 1);
    if (W > 1 && !G && !p2p) PMFS_NCCL(A.GroupStart());
    for (int k = 0; k < W; k++) {
        // peers in rotated order: in "phase" k every rank writes to a different owner (rank + k), so no owner's NVLink
        // ingress has all the ranks' slices queued up on it at once
        const int p = (me + k) % W;
        const unsigned long long sc = my_cuts[p + 1] - my_cuts[p];            // what I send to p
        const unsigned long long rcn = all[p * (W + 1) + me + 1] - all[p * (W + 1) + me];  // what p sends to me
        const int32_t own_p = m->parts.bound[p + 1] - m->parts.bound[p];
        if (p == me) {
            if (sc) PMFS_CUDA(cudaMemcpyAsync(m->d_inbox.ptr + sp.roff[me], mine + my_cuts[me], sc * sizeof(unsigned long long),
                                              cudaMemcpyDeviceToDevice, c->stream));
            PMFS_CUDA(cudaMemcpyAsync(m->d_noff_in.ptr + (size_t)me * (n_own + 1), c->d_noff.ptr + m->parts.bound[me],
                                      ((size_t)n_own + 1) * sizeof(uint32_t), cudaMemcpyDeviceToDevice, c->stream));
            continue;
        }
        if (G) {  // pull from the source's buffers
            if (rcn) PMFS_CUDA(cudaMemcpyAsync(m->d_inbox.ptr + sp.roff[p], G->skeys[p] + all[p * (W + 1) + me], rcn * sizeof(unsigned long long),
                                               cudaMemcpyDefault, c->stream));
            PMFS_CUDA(cudaMemcpyAsync(m->d_noff_in.ptr + (size_t)p * (n_own + 1), G->noff[p] + m->parts.bound[me],
                                      ((size_t)n_own + 1) * sizeof(uint32_t), cudaMemcpyDefault, c->stream));
            continue;
        }
        if (p2p) {  // push: my slice for p goes to the place in p's inbox that the cut points give it
            unsigned long long at = 0;
            for (int s2 = 0; s2 < me; s2++) at += all[s2 * (W + 1) + p + 1] - all[s2 * (W + 1) + p];
            if (sc) PMFS_CUDA(cudaMemcpyAsync(m->peer_inbox[p] + at, mine + my_cuts[p], sc * sizeof(unsigned long long), cudaMemcpyDefault, c->stream));
            PMFS_CUDA(cudaMemcpyAsync(m->peer_noff_in[p] + (size_t)me * ((size_t)own_p + 1), c->d_noff.ptr + m->parts.bound[p],
                                      ((size_t)own_p + 1) * sizeof(uint32_t), cudaMemcpyDefault, c->stream));
            continue;
        }
        if (sc) PMFS_NCCL(A.Send(mine + my_cuts[p], (size_t)sc, ncclUint64, p, m->comm, c->stream));
        if (rcn) PMFS_NCCL(A.Recv(m->d_inbox.ptr + sp.roff[p], (size_t)rcn, ncclUint64, p, m->comm, c->stream));
        PMFS_NCCL(A.Send(c->d_noff.ptr + m->parts.bound[p], (size_t)own_p + 1, ncclUint32, p, m->comm, c->stream));
        PMFS_NCCL(A.Recv(m->d_noff_in.ptr + (size_t)p * (n_own + 1), (size_t)n_own + 1, ncclUint32, p, m->comm, c->stream));
    }
    if (W > 1 && !G && !p2p) PMFS_NCCL(A.GroupEnd());
    if (p2p && (rc = comm_barrier(m, false))) return rc;  // everybody's copies into everybody's inbox are complete
    PMFS_CUDA(cudaEventRecord(m->ev[1], c->stream));
    // 4. stitch
    int pos_bits = 1;
    while ((1ll << pos_bits) < total_cols) pos_bits++;
    sp.inbox = m->d_inbox.ptr;
    sp.noff_in = m->d_noff_in.ptr;
    sp.world = W;
    sp.n_own = n_own;
    sp.node_base = m->parts.bound[me];
    sp.pos_bits = pos_bits;
    if (n_own > 0) {
        stitch_count_kernel<<<(unsigned)((n_own + 255) / 256), 256, 0, c->stream>>>(sp, m->d_total.ptr);
        size_t tmp = 0;
        PMFS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp, m->d_total.ptr, m->d_start.ptr, n_own, c->stream));
        if ((rc = m->d_scan_tmp.ensure(tmp + 16))) return rc;
        PMFS_CUDA(cub::DeviceScan::ExclusiveSum(m->d_scan_tmp.ptr, tmp, m->d_total.ptr, m->d_start.ptr, n_own, c->stream));
        if (c->n_sms == 0) PMFS_CUDA(cudaDeviceGetAttribute(&c->n_sms, cudaDevAttrMultiProcessorCount, c->device));
        const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>((int64_t)c->n_sms * 16, ((int64_t)n_own + 7) / 8));
        stitch_copy_kernel<<<grid, 256, 0, c->stream>>>(sp, m->d_start.ptr, c->d_out.ptr);
        PMFS_CUDA(cudaGetLastError());
        c->timing.n_launches += 2;
    }
    PMFS_CUDA(cudaEventRecord(m->ev[2], c->stream));
    PMFS_CUDA(cudaStreamSynchronize(c->stream));
    if (G) G->bar.wait();  // nobody may start its next pass (and overwrite its records) before everybody has pulled
    cudaEventElapsedTime(&m->exchange_ms, m->ev[0], m->ev[1]);
    cudaEventElapsedTime(&m->stitch_ms, m->ev[1], m->ev[2]);
    c->timing.exchange_ms = m->exchange_ms;
    c->timing.stitch_ms = m->stitch_ms;
    c->timing.total_ms += m->exchange_ms + m->stitch_ms;
    c->last_n_out = (int64_t)total;
    *n_out = (int64_t)total;
    if (d_out) *d_out = c->d_out.ptr;
    return PMFS_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// One process, several GPUs: what a C++ panmanUtils links against.  pmfs_multi owns one context and one host thread per
// device and an NCCL communicator over them (ncclCommInitAll); pmfs_multi_run takes the alignment in host memory, cuts
// it into column windows that fit, gives every device a contiguous share of each window (uploaded by the device's own
// thread), gathers by node range and returns ONE (node,pos)-ordered host array -- the contract of pmfs_run.
// ---------------------------------------------------------------------------------------------------------------------
struct pmfs_multi {
    int n = 0;
    std::vector<pmfs_ctx*> ctx;
    std::vector<pmfs_comm*> comm;
    std::vector<pmfs_mut> result;
    std::string error;
};

extern "C" {

void pmfs_multi_destroy(pmfs_multi* m) {
    if (!m) return;
    for (auto* x : m->comm) pmfs_comm_destroy(x);
    for (auto* x : m->ctx) pmfs_destroy(x);
    delete m;
}

int pmfs_multi_create(pmfs_multi** out, const pmfs_tree* tree, int n_gpus, const int* devices) {
    if (!out || !tree || n_gpus < 1 || n_gpus > kMaxRanks) {
        g_error = "pmfs_multi_create: bad argument (1 to 16 GPUs)";
        return PMFS_ERR_ARG;
    }
    *out = nullptr;
    const char* tr0 = getenv("PMFS_MULTI_TRANSPORT");
    const bool copy_transport = tr0 && std::string(tr0) == "copy";
    NcclApi& A = nccl_api();
    if (n_gpus > 1 && !copy_transport && !A.err.empty()) {
        g_error = "pmfs_multi_create: " + A.err;
        return PMFS_ERR_CUDA;
    }
    pmfs_multi* m = new pmfs_multi();
    m->n = n_gpus;
    std::vector<int> devs(n_gpus);
    for (int i = 0; i < n_gpus; i++) devs[i] = devices ? devices[i] : i;
    for (int i = 0; i < n_gpus; i++) {
        pmfs_ctx* c = nullptr;
        int rc = pmfs_create(&c, tree, devs[i]);
        if (rc) {
            pmfs_multi_destroy(m);
            return rc;
        }
        m->ctx.push_back(c);
    }
    // PMFS_MULTI_TRANSPORT=copy: device-to-device copies between the contexts instead of NCCL (also lets several
    // "ranks" share one GPU, which is how this driver is tested on a single GPU)
    const char* tr = getenv("PMFS_MULTI_TRANSPORT");
    if (tr && std::string(tr) == "copy") {
        m->comm.assign((size_t)n_gpus, nullptr);
        int rc = pmfs_comm_create_local(m->comm.data(), m->ctx.data(), n_gpus);
        if (rc) {
            pmfs_multi_destroy(m);
            return rc;
        }
        *out = m;
        return PMFS_OK;
    }
    std::vector<ncclComm_t> comms(n_gpus, nullptr);
    if (n_gpus > 1) {
        ncclResult_t r = A.CommInitAll(comms.data(), n_gpus, devs.data());
        if (r != ncclSuccess) {
            g_error = std::string("ncclCommInitAll: ") + A.GetErrorString(r);
            pmfs_multi_destroy(m);
            return PMFS_ERR_CUDA;
        }
    }
    for (int i = 0; i < n_gpus; i++) {
        pmfs_comm* x = new pmfs_comm();
        x->ctx = m->ctx[i];
        x->comm = comms[i];
        x->rank = i;
        x->world = n_gpus;
        m->comm.push_back(x);
        int rc = comm_finish_create(x);
        if (rc) {
            pmfs_multi_destroy(m);
            return rc;
        }
    }
    *out = m;
    return PMFS_OK;
}

int pmfs_multi_run(pmfs_multi* m, int algo, const uint8_t* msa, int64_t row_stride, int32_t n_rows, int64_t n_cols,
                   const uint8_t* cons, const uint8_t* forced, int64_t window_cols, pmfs_mut** out, int64_t* n_out) {
    if (!m || !msa || !cons || !out || !n_out || n_cols <= 0 || n_rows < 0 || row_stride * 2 < n_cols || window_cols < 0 ||
        (algo != PMFS_FITCH && algo != PMFS_SANKOFF)) {
        g_error = "pmfs_multi_run: bad argument";
        return PMFS_ERR_ARG;
    }
    const int N = m->n;
    const int64_t G_all = (n_cols + 31) / 32, used = (n_cols + 1) / 2;
    std::vector<int> rcs(N, PMFS_OK);
    std::vector<std::string> errs(N);
    // per window and rank: the owner's records, downloaded by the owner's thread
    std::vector<std::vector<std::vector<pmfs_mut>>> got;  // [window][rank]
    std::vector<int64_t> win_g0;                          // first column group of every window
    int64_t gw = 0;                                       // column groups per rank and window (set by rank 0 below)
    Barrier bar(N);
    auto worker = [&](int r) {
        pmfs_ctx* c = m->ctx[r];
        pmfs_comm* x = m->comm[r];
        auto fail = [&](int rc) {
            rcs[r] = rc;
            errs[r] = g_error;
        };
        cudaSetDevice(c->device);
        // window size: what fits next to one staged window share on every device
        int64_t per_rank = (G_all + N - 1) / N, bc = 0;
        if (window_cols > 0) per_rank = std::max<int64_t>(1, std::min(per_rank, (window_cols / N + 31) / 32));
        int rc = pmfs_comm_batch_cols(x, algo, per_rank * 32, ((int64_t)n_rows + 1) / 2 + 2, &bc);  // + the staged window
        if (rc) fail(rc);
        if (r == 0) {
            gw = std::max<int64_t>(1, (bc + 31) / 32);
            for (int64_t g0 = 0; g0 < G_all; g0 += gw * N) win_g0.push_back(g0);
            got.assign(win_g0.size(), std::vector<std::vector<pmfs_mut>>(N));
        }
        bar.wait();
        bool any_fail = false;
        for (int i = 0; i < N; i++) any_fail |= rcs[i] != PMFS_OK;
        if (any_fail) return;
        const int64_t bstride = gw * 16;
        if ((rc = c->d_msa.ensure((size_t)std::max<int32_t>(1, n_rows) * (size_t)bstride))) fail(rc);
        if (!rc && (rc = c->d_cons.ensure((size_t)gw * 32))) fail(rc);
        if (!rc && forced && (rc = c->d_forced.ensure((size_t)gw * 32))) fail(rc);
        for (size_t w = 0; w < win_g0.size(); w++) {
            // this rank's share of the window: groups [g0, g1) (possibly empty at the ragged end)
            const int64_t g0 = std::min(G_all, win_g0[w] + r * gw), g1 = std::min(G_all, g0 + gw);
            const int64_t c0 = std::min(n_cols, g0 * 32), nc = std::max<int64_t>(0, std::min(n_cols, g1 * 32) - c0);
            int64_t n = 0;
            pmfs_mut* d_out = nullptr;
            if (rcs[r] == PMFS_OK) {
                cudaError_t e = cudaSuccess;
                if (nc > 0) {
                    const int64_t wbytes = std::min<int64_t>(bstride, used - g0 * 16);
                    if (wbytes < bstride) e = cudaMemsetAsync(c->d_msa.ptr, 0, (size_t)std::max<int32_t>(1, n_rows) * bstride, c->stream);
                    if (e == cudaSuccess && n_rows > 0)
                        e = cudaMemcpy2DAsync(c->d_msa.ptr, (size_t)bstride, msa + g0 * 16, (size_t)row_stride, (size_t)wbytes, (size_t)n_rows,
                                              cudaMemcpyHostToDevice, c->stream);
                    if (e == cudaSuccess) e = cudaMemcpyAsync(c->d_cons.ptr, cons + c0, (size_t)nc, cudaMemcpyHostToDevice, c->stream);
                    if (e == cudaSuccess && forced) e = cudaMemcpyAsync(c->d_forced.ptr, forced + c0, (size_t)nc, cudaMemcpyHostToDevice, c->stream);
                }
                if (e != cudaSuccess) {
                    g_error = std::string("pmfs_multi_run upload: ") + cudaGetErrorString(e);
                    fail(PMFS_ERR_CUDA);
                }
            }
            // (a rank whose share is empty -- the ragged end of a narrow alignment -- still takes part in the exchange)
            if (rcs[r] == PMFS_OK) {
                rc = pmfs_comm_run(x, algo, c->d_msa.ptr, bstride, n_rows, nc, c0, n_cols, c->d_cons.ptr, forced ? c->d_forced.ptr : nullptr,
                                   &d_out, &n);
                if (rc) fail(rc);
            }
            if (rcs[r] == PMFS_OK && n > 0) {
                got[w][r].resize((size_t)n);
                cudaError_t e = cudaMemcpyAsync(got[w][r].data(), d_out, (size_t)n * sizeof(pmfs_mut), cudaMemcpyDeviceToHost, c->stream);
                if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
                if (e != cudaSuccess) {
                    g_error = std::string("pmfs_multi_run download: ") + cudaGetErrorString(e);
                    fail(PMFS_ERR_CUDA);
                }
            }
            if (rcs[r] != PMFS_OK) return;  // (the other ranks then fail in NCCL or hang: see the note in pmfs.h)
        }
    };
    std::vector<std::thread> th;
    for (int r = 0; r < N; r++) th.emplace_back(worker, r);
    for (auto& t : th) t.join();
    for (int r = 0; r < N; r++)
        if (rcs[r] != PMFS_OK) {
            g_error = "pmfs_multi_run (GPU " + std::to_string(r) + "): " + errs[r];
            return rcs[r];
        }
    // windows are column-ordered and, inside a window, ranks are node-ordered: stitch the windows by node
    const size_t V = (size_t)m->ctx[0]->plan.n_nodes;
    std::vector<int64_t> at(V + 1, 0);
    size_t total = 0;
    for (auto& w : got)
        for (auto& part : w) {
            total += part.size();
            for (auto& rec : part) at[(size_t)rec.node + 1]++;
        }
    for (size_t v = 0; v < V; v++) at[v + 1] += at[v];
    m->result.resize(total);
    for (auto& w : got)
        for (auto& part : w) {
            for (auto& rec : part) m->result[(size_t)at[(size_t)rec.node]++] = rec;
            std::vector<pmfs_mut>().swap(part);
        }
    *out = m->result.data();
    *n_out = (int64_t)total;
    return PMFS_OK;
}

}  // extern "C"
