// The whole level loop as ONE cooperative kernel: per level the histogram pass (with the fused repartition),
// a grid-wide barrier, the scan of every frontier node, a barrier, the planning of the next level plus the
// clearing of its histogram region, a barrier -- and after the last level the final repartition.  Replaces two
// launches (and their launch gaps and per-launch start-up) per level of the stream-ordered loop; the phases are
// the same device functions (hist_block, scan_node, plan_level).  Used when every level's histogram fits the
// shared-memory path; included by avt_kernels.cuh.
#pragma once
#include <cooperative_groups.h>

namespace avt {

namespace cg = cooperative_groups;

struct LevelParams {
    HistArgs ha;                 // histogram pass of this level (ha.nid -> ha.nid_out when `repart`)
    PlanArgs pa;                 // planning of the next level (zero_ptr unused: the grid clears cooperatively)
    const uint32_t* Hc;          // complete counts the scan reads (ha.H on one GPU, the local sums with the peer exchange)
    const int32_t* n_nodes_dev;  // true frontier size of this level
    uint32_t* zero_next_ptr;     // next level's histogram region
    long long zero_next_words;
    int n_chunks, n_groups, repart;
};

struct PersistArgs {
    const LevelParams* levels;
    int n_levels;
    SplitTable st;
    const uint8_t* order;
    int criterion;
    double* av;
    avt_node_best* best;
    avt_split_eval* best_eval;
    // final repartition: apply the last planned split to the node ids
    uint16_t* final_nid;
    const void* const* cols;
    const int32_t* wide;
    const int32_t* final_node_base;
    const int32_t* node_attr;
    const uint16_t* child_lut;
    int lut_w, T;
    long long N, Npad;
    unsigned long long* phase_ns;   // optional [n_levels*3 + 2] %globaltimer stamps written by block 0 (developer timing)
};

// rows [base, base+8) of tree t: nid <- child_lut[node][code of the node's split attribute]
__device__ __forceinline__ void repartition_rows8(uint16_t* nid, const void* const* cols, const int32_t* wide, int nb,
                                                  const int32_t* node_attr, const uint16_t* child_lut, int lut_w, long long base, long long N) {
    uint4 v = *reinterpret_cast<const uint4*>(nid + base);
    uint32_t wds[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        uint32_t old = (wds[j >> 1] >> ((j & 1) * 16)) & 0xFFFFu;
        uint32_t nw = AVT_NODE_NONE;
        if (old != AVT_NODE_NONE && base + j < N) {
            const int g = nb + (int)old;
            const int at = node_attr[g];
            if (at >= 0) {
                const long long r = base + j;
                uint32_t code = wide[at] ? ((const uint16_t*)cols[at])[r] : ((const uint8_t*)cols[at])[r];
                nw = child_lut[(long long)g * lut_w + code];
            }
        }
        wds[j >> 1] = (wds[j >> 1] & ~(0xFFFFu << ((j & 1) * 16))) | (nw << ((j & 1) * 16));
    }
    *reinterpret_cast<uint4*>(nid + base) = make_uint4(wds[0], wds[1], wds[2], wds[3]);
}

template <bool HAS_W, bool WIDE>
__global__ void __launch_bounds__(kScanThreads) grow_persistent_kernel(const PersistArgs g) {
    extern __shared__ uint32_t sh[];
    __shared__ LevelParams lp;
    __shared__ int s_warp_tot[kScanWarps];
    __shared__ int s_running;
    cg::grid_group grid = cg::this_grid();
    auto stamp = [&](int slot) {
        if (g.phase_ns && blockIdx.x == 0 && threadIdx.x == 0) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
            g.phase_ns[slot] = t;
        }
    };
    stamp(0);
    for (int level = 0; level < g.n_levels; ++level) {
        __syncthreads();
        {   // this level's parameters, once per block
            const uint32_t* src = reinterpret_cast<const uint32_t*>(g.levels + level);
            uint32_t* dst = reinterpret_cast<uint32_t*>(&lp);
            for (int i = threadIdx.x; i < (int)(sizeof(LevelParams) / 4); i += blockDim.x) dst[i] = src[i];
        }
        __syncthreads();
        // ---- (a) histogram pass: work items = (tree, row chunk, attribute group), tree fastest
        const int work = g.T * lp.n_chunks * lp.n_groups;
        for (int w = blockIdx.x; w < work; w += gridDim.x) {
            const int t = w % g.T, chunk = (w / g.T) % lp.n_chunks, grp = w / (g.T * lp.n_chunks);
            __syncthreads();
            for (int i = threadIdx.x; i < lp.ha.zero_words; i += blockDim.x) sh[i] = 0;
            __syncthreads();
            if (lp.repart) hist_block<HAS_W, WIDE, true>(lp.ha, sh, t, chunk, lp.n_chunks, grp, lp.n_groups, (uint32_t)work);
            else hist_block<HAS_W, WIDE, false>(lp.ha, sh, t, chunk, lp.n_chunks, grp, lp.n_groups, (uint32_t)work);
        }
        __threadfence();
        grid.sync();
        stamp(1 + level * 3);
        // ---- (b) scan: one node per block at a time
        const int NF = *lp.n_nodes_dev;
        for (int node = blockIdx.x; node < NF; node += gridDim.x) {
            __syncthreads();
            scan_node(lp.Hc, g.st, nullptr, g.order, g.criterion, g.av, g.best, g.best_eval, node, lp.pa);
        }
        __threadfence();
        grid.sync();
        stamp(2 + level * 3);
        // ---- planning of the next level (one block) + clearing of its histogram region (everyone)
        if (blockIdx.x == 0) plan_level(lp.pa, g.st, s_warp_tot, &s_running);
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < lp.zero_next_words; i += (long long)gridDim.x * blockDim.x)
            lp.zero_next_ptr[i] = 0;
        __threadfence();
        grid.sync();
        stamp(3 + level * 3);
    }
    // ---- final repartition
    for (int t = 0; t < g.T; ++t) {
        uint16_t* nid = g.final_nid + (long long)t * g.Npad;
        const int nb = g.final_node_base[t];
        for (long long base = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8; base < g.N; base += (long long)gridDim.x * blockDim.x * 8)
            repartition_rows8(nid, g.cols, g.wide, nb, g.node_attr, g.child_lut, g.lut_w, base, g.N);
    }
    grid.sync();
    stamp(1 + g.n_levels * 3);
}

}  // namespace avt
