// One translation unit per tile shape: PMFS_WALK_INSTANCE(NW, NG) instantiates the walk kernels of that shape and
// defines the table entry pmfs.cu reaches them through (walk_params.hpp).
#pragma once

#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "walk_kernels.cuh"

namespace pmfs {

#ifdef PMFS_WAITPROF
static void prof_reset_impl() {
    unsigned long long z[16] = {0};
    cudaMemcpyToSymbol(g_waitprof, z, sizeof(z));
    cudaMemcpyToSymbol(g_trace_n, z, sizeof(unsigned int));
}
static void prof_dump_impl(bool fwd, unsigned grid, int per_sm, int n_items) {
    unsigned long long z[16];
    cudaMemcpyFromSymbol(z, g_waitprof, sizeof(z));
    const double W = (double)z[5], Pr = (double)z[10];
    fprintf(stderr, "[waitprof %s grid %u per_sm %d items %d] walkers: desc %.1f%% prog %.1f%% leaf %.1f%% set %.1f%% barsync %.1f%% | "
            "producer: dep %.1f%% desc_empty %.1f%% leaf_empty %.1f%% set_empty %.1f%% | walker warp-cycles %.3g\n",
            fwd ? "fwd" : "bwd", grid, per_sm, n_items, 100 * z[0] / W, 100 * z[1] / W, 100 * z[2] / W, 100 * z[3] / W, 100 * z[4] / W,
            100 * z[6] / Pr, 100 * z[7] / Pr, 100 * z[8] / Pr, 100 * z[9] / Pr, W);
    if (const char* dir = getenv("PMFS_TRACE_DIR")) {
        unsigned int tn = 0;
        cudaMemcpyFromSymbol(&tn, g_trace_n, sizeof(tn));
        tn = tn < (unsigned)kTraceCap ? tn : (unsigned)kTraceCap;
        std::vector<ulonglong4> tr(tn);
        if (tn) cudaMemcpyFromSymbol(tr.data(), g_trace, (size_t)tn * sizeof(ulonglong4));
        const std::string path = std::string(dir) + (fwd ? "/trace_fwd.bin" : "/trace_bwd.bin");
        if (FILE* fp = fopen(path.c_str(), "wb")) {
            fwrite(tr.data(), sizeof(ulonglong4), tn, fp);
            fclose(fp);
        }
    }
}
#endif

template <class CFG>
static WalkKernelSet make_kernel_set() {
    WalkKernelSet s = {};
    s.shape.nw = CFG::NW;
    s.shape.ng = CFG::NG;
    s.fwd[kAlgoFitch][0] = (const void*)fwd_kernel<CFG, kAlgoFitch, false>;
    s.fwd[kAlgoFitch][1] = (const void*)fwd_kernel<CFG, kAlgoFitch, true>;
    s.fwd[kAlgoSankoff][0] = (const void*)fwd_kernel<CFG, kAlgoSankoff, false>;
    s.fwd[kAlgoSankoff][1] = (const void*)fwd_kernel<CFG, kAlgoSankoff, true>;
    s.bwd[kAlgoFitch] = (const void*)bwd_kernel<CFG, kAlgoFitch>;
    s.bwd[kAlgoSankoff] = (const void*)bwd_kernel<CFG, kAlgoSankoff>;
    s.smem_fixed = kSmemLeaf;
    s.leaf_unit_bytes = CFG::UnitBytes;
    s.set_slot_bytes = CFG::SetBytes;
    s.fwd_smem_per_level = (size_t)CFG::NG * 4 * CFG::TPB * sizeof(uint4);
    s.bwd_smem_per_level = (size_t)CFG::NG * 5 * CFG::TPB * sizeof(uint32_t);
#ifdef PMFS_WAITPROF
    s.prof_reset = prof_reset_impl;
    s.prof_dump = prof_dump_impl;
#endif
    return s;
}

}  // namespace pmfs

#define PMFS_WALK_INSTANCE(NW, NG)                                                     \
    namespace pmfs {                                                                   \
    const WalkKernelSet* walk_kernel_set_##NW##x##NG() {                               \
        static const WalkKernelSet s = make_kernel_set<TileCfg<NW, NG>>();             \
        return &s;                                                                     \
    }                                                                                  \
    }
