// Hand-written sm_100a kernels of the split-evaluation path.
//   (a) hist_smem_kernel / hist_global_kernel : (tree,node,attr,bin,class) count histograms
//   (b) scan_kernel / eval_kernel / cpg_kernel: fused fp64 impurity scan + per-node arg-min
//       scan_fast_kernel (avt_scan_fast.cuh)  : latency-optimised scan + level planning for small split tables
//   (c) repartition_kernel / repartition_fast_kernel : row -> child node
//       csv_* kernels (avt_csv.cuh)           : CSV text -> bin codes (the step before the path)
//       predict_kernel (avt_predict.cuh)      : scoring with the emitted models
// All of it is HBM-/atomic-bound integer work: no tensor cores (DESIGN.md "Roofline").
#pragma once
#include "avt_device.cuh"

namespace avt {

constexpr int kHistThreads = 512;
constexpr int kRowsPerThread = 16;  // one 128-bit load of uint8 codes
constexpr int kScanThreads = 256;
constexpr int kScanWarps = kScanThreads / 32;
constexpr int kScanWordsPerWarp = 1024;  // per-warp prefix staging (4 KB; static smem <= 48 KB)

// ------------------------------------------------------------------------------------------------
// (a) histogram
// ------------------------------------------------------------------------------------------------
struct HistArgs {
    const void* const* cols;   // [A] device column pointers (uint8 or uint16 codes, padded to 16 rows)
    const uint8_t* cls;        // [Npad]
    const uint16_t* nid;       // [T][Npad] node ids read by this pass
    const uint8_t* wts;        // [T][Npad] or nullptr
    const uint8_t* has_w;      // [T] (nullptr when wts == nullptr)
    uint32_t* H;               // [NF][TB][C]
    const int32_t* node_base;  // [T+1] frontier being counted
    // attribute groups (one block handles one group): entry e in [grp_start[g], grp_start[g+1])
    const int32_t* ent_attr;    // attribute index
    const int32_t* ent_locoff;  // first bin of the attribute inside its group
    const int32_t* ent_globoff; // first bin of the attribute in the full bin space
    const int32_t* ent_wide;    // 1 if the column holds uint16 codes
    const int32_t* grp_start;   // [G+1]
    const int32_t* grp_bins;    // [G] bins in the group
    const int32_t* grp_rlog;    // [G] log2(replicas)
    const int32_t* grp_l2g;     // concatenated local-bin -> global-bin maps
    const int32_t* grp_l2g_off; // [G]
    long long N, Npad;
    int C, TB, T;
    // fused repartition (kernel (c) folded into kernel (a)): with nid_out != nullptr every row is first
    // moved from its node of the previous frontier to the child its bin selects, the new id is stored
    // to nid_out and counted.  `nid` and `nid_out` are different buffers (blocks of different attribute
    // groups redo the same move).
    uint16_t* nid_out;
    const int32_t* prev_node_base;  // [T+1]
    const int32_t* node_attr;       // [NF_prev] split attribute of each previous node, -1 = not expanded
    const uint16_t* child_lut;      // [NF_prev][lut_w]
    const int32_t* wide;            // [A]
    int lut_w;
    int codes_in_regs;  // 1: one group, <= 4 uint8 attributes, entry q == attribute q (the split code is in registers)
    // multi-GPU peer exchange: the last block of the pass to finish tells every rank that this rank's
    // partial histogram is complete (no separate signalling launch)
    int zero_words;                 // shared-memory words the block may clear before it waits for the previous launch
    const uint8_t* skip;            // [NF] nodes of this frontier whose rows are not counted (sibling subtraction), nullptr = none
    uint32_t* const* signal_flags;  // [signal_ranks] this rank's flag word inside every rank's memory (nullptr: no signalling)
    uint32_t* signal_ticket;        // zero-initialised block counter
    uint32_t signal_epoch;
    int signal_ranks;
    // bit-packed codes: when the data set has <= 4 narrow attributes whose codes fit one byte together, the pass reads that
    // one byte per row instead of one byte per attribute (pack_codes_kernel builds it when the data set is loaded)
    int32_t* bad;            // hist_global_kernel only, nullable: rows whose node id, bin code or class code is outside the declared
                             // shape are not counted and *bad is set to 1 (callers that stage unvalidated data: avt_histogram_static)
    const uint8_t* packed;   // [Npad] or nullptr
    uint32_t pk_shift[4];    // bit position of attribute q's code
    uint32_t pk_mask4[4];    // its mask, replicated into the four bytes of a word
    int packed_bins;    // 1: every attribute group has <= 256 bins (group-local bin indices fit a byte)
    int repart_smem;    // 1: codes_in_regs and the previous level's tables fit the block's shared memory (fast path)
};

// Shared-memory copy of the previous level's tables for the fast fused repartition.
struct RepartSmem {
    const uint32_t* sel;   // [n_prev+1] byte-permute selector 0x4440|attr of each previous node
    const uint16_t* lut;   // [n_prev+1][lut_w] bin -> child (entry n_prev: all NONE)
    uint32_t n_prev;
    bool fast;
    const uint8_t* skip;   // [n_nodes+1] 1 = node is derived by subtraction, its rows are not counted (nullptr: count all)
};

// Fast fused repartition of 16 rows (<= 4 uint8 attributes, previous frontier's tables in shared memory): new node id
// of every row into nw[] (two ids per word).  v[] holds the rows' codes, one vector per attribute.
__device__ __forceinline__ void repart16_fast(const uint4& vn0, const uint4& vn1, uint4 (&v)[4], int n0, const RepartSmem& rp, int lut_w,
                                              uint32_t (&nw)[8]) {
    if (n0 < 4) v[3] = make_uint4(0, 0, 0, 0);
    if (n0 < 3) v[2] = make_uint4(0, 0, 0, 0);
    if (n0 < 2) v[1] = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int w = 0; w < 4; ++w) {
        const uint32_t a0 = u4_word(v[0], w), a1 = u4_word(v[1], w), a2 = u4_word(v[2], w), a3 = u4_word(v[3], w);
        const uint32_t t0 = __byte_perm(a0, a1, 0x5140), t1 = __byte_perm(a2, a3, 0x5140);
        const uint32_t t2 = __byte_perm(a0, a1, 0x7362), t3 = __byte_perm(a2, a3, 0x7362);
        const uint32_t pk[4] = {__byte_perm(t0, t1, 0x5410), __byte_perm(t0, t1, 0x7632), __byte_perm(t2, t3, 0x5410),
                                __byte_perm(t2, t3, 0x7632)};
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int j = w * 4 + r;
            const uint32_t old = j < 8 ? u4_half(vn0, j) : u4_half(vn1, j - 8);
            const uint32_t idx = min(old, rp.n_prev);  // 0xFFFF -> the phantom entry (all-NONE row)
            const uint32_t code = __byte_perm(pk[r], 0u, rp.sel[idx]);
            const uint32_t nn = rp.lut[idx * lut_w + code];
            if (j & 1) nw[j >> 1] |= nn << 16;
            else nw[j >> 1] = nn;
        }
    }
}

// hist_rows16: 16 consecutive rows of one tree for one attribute group.
//   rowkey[j] = ((node * GB * C + cls) << rlog) + replica   (word index of the row's class slot of bin 0)
//   counter   = rowkey[j] + ((locoff * C) << rlog) + code * (C << rlog)          (all kept as byte addresses)
template <bool HAS_W, bool WIDE, bool REPART>
__device__ __forceinline__ void hist_rows16(uint32_t sh_base, const HistArgs& a, int e0, int ne, long long base,
                                            const uint16_t* nid, uint16_t* nid_out, const uint8_t* wt,
                                            uint32_t node_stride, uint32_t cs4, int rlog, uint32_t rep, int prev_nb,
                                            uint32_t phantom, const RepartSmem& rp) {
    uint4 vn0 = ld_stream(nid + base);
    uint4 vn1 = ld_stream(nid + base + 8);
    const uint4 vc = ld_stream(a.cls + base);
    uint4 vw = make_uint4(0, 0, 0, 0);
    if (HAS_W) vw = ld_stream(wt + base);
    uint4 v[4];
    const int n0 = ne < 4 ? ne : 4;
    if (!WIDE) {
        if (a.packed) {
            // entry q == attribute q here (codes_in_regs); unpack 4 rows at a time: shift, then mask every byte
            const uint4 vp = ld_stream(a.packed + base);
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (q < n0) {
                    const uint32_t sh = a.pk_shift[q], m = a.pk_mask4[q];
                    v[q] = make_uint4((vp.x >> sh) & m, (vp.y >> sh) & m, (vp.z >> sh) & m, (vp.w >> sh) & m);
                }
        } else {
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (q < n0) v[q] = ld_stream((const char*)a.cols[a.ent_attr[e0 + q]] + base);
        }
    }
    if (REPART) {
        uint32_t nw[8];
        if (!WIDE && rp.fast) {
            // Fast form (<= 4 uint8 attributes, small previous frontier): the split attribute (as a byte-
            // permute selector) and the bin -> child rows of every previous node sit in shared memory; the
            // row's four codes are gathered into one word by 4x4 byte transposes, so choosing "the code
            // of the node's split attribute" is a single PRMT.
            repart16_fast(vn0, vn1, v, n0, rp, a.lut_w, nw);
        } else {
#pragma unroll
            for (int j = 0; j < kRowsPerThread; ++j) {
                const uint32_t old = j < 8 ? u4_half(vn0, j) : u4_half(vn1, j - 8);
                uint32_t nn = AVT_NODE_NONE;
                if (old != AVT_NODE_NONE) {
                    const int gi = prev_nb + (int)old;
                    const int at = __ldg(a.node_attr + gi);
                    if (at >= 0) {
                        const uint32_t code = a.wide[at] ? ((const uint16_t*)a.cols[at])[base + j] : ((const uint8_t*)a.cols[at])[base + j];
                        nn = __ldg(a.child_lut + (long long)gi * a.lut_w + code);
                    }
                }
                if (j & 1) nw[j >> 1] |= nn << 16;
                else nw[j >> 1] = nn;
            }
        }
        vn0 = make_uint4(nw[0], nw[1], nw[2], nw[3]);
        vn1 = make_uint4(nw[4], nw[5], nw[6], nw[7]);
        *reinterpret_cast<uint4*>(nid_out + base) = vn0;
        *reinterpret_cast<uint4*>(nid_out + base + 8) = vn1;
    }
    // byte address (shared window) of the row's class slot of bin 0; rows outside the frontier (and rows
    // with multiplicity 0) are sent to a phantom node behind the last real one, so the update loop has
    // no branches; the phantom counters are never flushed.
    uint32_t rowkey[kRowsPerThread];
#pragma unroll
    for (int j = 0; j < kRowsPerThread; ++j) {
        uint32_t node = j < 8 ? u4_half(vn0, j) : u4_half(vn1, j - 8);
        const bool dead = node == AVT_NODE_NONE || (HAS_W && u4_byte(vw, j) == 0);
        rowkey[j] = dead ? phantom : node;
    }
    if (rp.skip) {   // sibling subtraction: rows of derived nodes are not counted (tested once per 16 rows, not per row)
#pragma unroll
        for (int j = 0; j < kRowsPerThread; ++j) rowkey[j] = rp.skip[rowkey[j]] ? phantom : rowkey[j];
    }
    {
        const uint32_t ns4 = node_stride * 4u, cl4 = 4u << rlog, base_rep = sh_base + rep * 4u;
#pragma unroll
        for (int j = 0; j < kRowsPerThread; ++j) rowkey[j] = rowkey[j] * ns4 + (u4_byte(vc, j) * cl4 + base_rep);   // two IMADs
    }
    if (!WIDE) {
        for (int c0 = 0; c0 < ne; c0 += 4) {
            const int nq = ne - c0 < 4 ? ne - c0 : 4;
            if (c0) {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (q < nq) v[q] = ld_stream((const char*)a.cols[a.ent_attr[e0 + c0 + q]] + base);
            }
            if (a.packed_bins) {
                // the group has <= 256 bins: the attribute's first bin is added to its 16 codes four at a time (no carry
                // between the bytes), which leaves extract + multiply-add + update per counter
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    if (q < nq) {
                        const uint32_t lo4 = (uint32_t)a.ent_locoff[e0 + c0 + q] * 0x01010101u;
                        const uint4 vv = make_uint4(v[q].x + lo4, v[q].y + lo4, v[q].z + lo4, v[q].w + lo4);
#pragma unroll
                        for (int j = 0; j < kRowsPerThread; ++j)
                            red_shared_add(rowkey[j] + u4_byte(vv, j) * cs4, HAS_W ? u4_byte(vw, j) : 1u);
                    }
                }
            } else {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    if (q < nq) {
                        const uint32_t offs = ((uint32_t)(a.ent_locoff[e0 + c0 + q] * a.C) << rlog) * 4u;
#pragma unroll
                        for (int j = 0; j < kRowsPerThread; ++j)
                            red_shared_add(rowkey[j] + offs + u4_byte(v[q], j) * cs4, HAS_W ? u4_byte(vw, j) : 1u);
                    }
                }
            }
        }
    } else {
        for (int e = e0; e < e0 + ne; ++e) {
            const char* col = (const char*)a.cols[a.ent_attr[e]];
            const bool wd = a.ent_wide[e] != 0;
            uint4 va, vb;
            if (wd) { va = ld_stream(col + base * 2); vb = ld_stream(col + base * 2 + 16); }
            else { va = ld_stream(col + base); vb = va; }
            const uint32_t offs = ((uint32_t)(a.ent_locoff[e] * a.C) << rlog) * 4u;
#pragma unroll
            for (int j = 0; j < kRowsPerThread; ++j) {
                const uint32_t code = wd ? (j < 8 ? u4_half(va, j) : u4_half(vb, j - 8)) : u4_byte(va, j);
                red_shared_add(rowkey[j] + offs + code * cs4, HAS_W ? u4_byte(vw, j) : 1u);
            }
        }
    }
}

// Shared-memory privatised histogram, optionally with the previous level's repartition folded in.
// grid = (T, chunks, G).  Block (t, chunk, g) counts the rows of its grid-strided 16-row vectors into a
// private copy of H[t][*][bins of group g][*] kept in shared memory, replicated 2^rlog times (copy =
// lane mod replicas) so that lanes of a warp that hit the same counter land in different banks; the
// block then folds the copies and adds them to global H.
// Work of one block: tree t, row chunk `chunk` of `n_chunks`, attribute group g (of n_groups).  `sh` is the
// block's dynamic shared memory, its first a.zero_words words already cleared.
template <bool HAS_W, bool WIDE, bool REPART>
__device__ __forceinline__ void hist_block(const HistArgs& a, uint32_t* sh, int t, int chunk, int n_chunks, int g, int n_groups,
                                           uint32_t total_blocks) {
    const int nb = a.node_base[t], n_nodes = a.node_base[t + 1] - nb;
    const int GB = a.grp_bins[g], rlog = a.grp_rlog[g];
    const int keys = n_nodes * GB * a.C;
    const int words = (keys + GB * a.C) << rlog;  // + one phantom node that swallows rows outside the frontier
    for (int i = a.zero_words + threadIdx.x; i < words; i += blockDim.x) sh[i] = 0;   // (only if the bound was not given)
    RepartSmem rp{nullptr, nullptr, 0u, false, nullptr};
    uint8_t* smem_tail = reinterpret_cast<uint8_t*>(sh + words);
    if (REPART && a.repart_smem) {
        const int pnb = a.prev_node_base[t], n_prev = a.prev_node_base[t + 1] - pnb;
        uint32_t* sel = sh + words;
        uint16_t* lut = reinterpret_cast<uint16_t*>(sel + n_prev + 1);
        for (int i = threadIdx.x; i <= n_prev; i += blockDim.x) {
            const int at = i < n_prev ? a.node_attr[pnb + i] : -1;
            sel[i] = 0x4440u | (uint32_t)(at < 0 ? 0 : at);
        }
        for (int i = threadIdx.x; i < (n_prev + 1) * a.lut_w; i += blockDim.x)
            lut[i] = i < n_prev * a.lut_w ? a.child_lut[(long long)pnb * a.lut_w + i] : (uint16_t)AVT_NODE_NONE;
        rp = RepartSmem{sel, lut, (uint32_t)n_prev, true, nullptr};
        smem_tail = reinterpret_cast<uint8_t*>(lut + (n_prev + 1) * a.lut_w);
    }
    if (a.skip) {
        for (int i = threadIdx.x; i <= n_nodes; i += blockDim.x) smem_tail[i] = i < n_nodes ? a.skip[nb + i] : 0;
        rp.skip = smem_tail;
    }
    __syncthreads();
    const uint32_t sh_base = (uint32_t)__cvta_generic_to_shared(sh);
    const uint16_t* nid = a.nid + (long long)t * a.Npad;
    uint16_t* nid_out = REPART ? a.nid_out + (long long)t * a.Npad : nullptr;
    const bool tree_has_w = HAS_W && a.has_w[t];
    const uint8_t* wt = HAS_W ? a.wts + (long long)t * a.Npad : nullptr;
    const uint32_t rep = (threadIdx.x & 31) & ((1u << rlog) - 1);
    const uint32_t cs4 = ((uint32_t)a.C << rlog) * 4u;
    const uint32_t node_stride = (uint32_t)(GB * a.C) << rlog;
    const int e0 = a.grp_start[g], ne = a.grp_start[g + 1] - e0;
    const int prev_nb = REPART ? a.prev_node_base[t] : 0;
    const int lane = threadIdx.x & 31;
    const long long stride = (long long)n_chunks * blockDim.x * kRowsPerThread;
    for (long long base = ((long long)chunk * blockDim.x + threadIdx.x) * kRowsPerThread; base < a.N; base += stride) {
        if (tree_has_w) hist_rows16<true, WIDE, REPART>(sh_base, a, e0, ne, base, nid, nid_out, wt, node_stride, cs4, rlog, rep, prev_nb, (uint32_t)n_nodes, rp);
        else hist_rows16<false, WIDE, REPART>(sh_base, a, e0, ne, base, nid, nid_out, wt, node_stride, cs4, rlog, rep, prev_nb, (uint32_t)n_nodes, rp);
    }
    __syncthreads();
    // fold the copies and add them to H: one warp per node, lanes over the node's (bin, class) slots
    const int* l2g = a.grp_l2g + a.grp_l2g_off[g];
    const int R = 1 << rlog;
    const int GBC = GB * a.C;
    const int warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    if (n_groups == 1) {
        // one group holds every attribute: local and global bin indices coincide, so the block's counters map onto
        // H[nb..nb+n_nodes) word for word -- a flat loop without the per-node bookkeeping below
        uint32_t* Hb = a.H + (long long)nb * a.TB * a.C;
        for (int k = threadIdx.x; k < keys; k += blockDim.x) {
            uint32_t sum = 0;
            for (int r = 0; r < R; ++r) sum += sh[(k << rlog) + ((r + lane) & (R - 1))];
            if (sum) atomicAdd(Hb + k, sum);
        }
    } else
    for (int node = warp; node < n_nodes; node += n_warps) {
        uint32_t* Hn = a.H + (long long)(nb + node) * a.TB * a.C;
        int lb = lane / a.C, c = lane - lb * a.C;       // (bin, class) of slot kk = lane; advanced without divisions
        const int step_b = 32 / a.C, step_c = 32 - step_b * a.C;
        for (int kk = lane; kk < GBC; kk += 32) {
            const int k = node * GBC + kk;
            uint32_t sum = 0;
            for (int r = 0; r < R; ++r) sum += sh[(k << rlog) + ((r + lane) & (R - 1))];
            if (sum) atomicAdd(&Hn[l2g[lb] * a.C + c], sum);
            lb += step_b;
            c += step_c;
            if (c >= a.C) { c -= a.C; ++lb; }
        }
    }
    if (a.signal_flags) {
        __shared__ int s_last;
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) {
            const uint32_t total = total_blocks;
            s_last = atomicAdd(a.signal_ticket, 1u) == total - 1;
            if (s_last) *a.signal_ticket = 0;
        }
        __syncthreads();
        // one lane per rank, so the system-scope fences and the NVLink stores overlap instead of queueing
        if (s_last && (int)threadIdx.x < a.signal_ranks) {
            __threadfence_system();
            st_release_sys(a.signal_flags[threadIdx.x], a.signal_epoch);
        }
    }
}

// Shared-memory privatised histogram, optionally with the previous level's repartition folded in.
// grid = (T, chunks, G); see hist_block.
template <bool HAS_W, bool WIDE, bool REPART>
__global__ void __launch_bounds__(kHistThreads) hist_smem_kernel(const HistArgs a) {
    extern __shared__ uint32_t sh[];
    pdl_launch_dependents();
    // clear the block's histogram copies (sized by the host's bound) while the previous launch may still run
    for (int i = threadIdx.x; i < a.zero_words; i += blockDim.x) sh[i] = 0;
    pdl_wait();   // node_base, the repartition tables and the node ids come from the previous launches
    __syncthreads();  // the cleared region may extend past this frontier's counters into where the tables are staged below
    hist_block<HAS_W, WIDE, REPART>(a, sh, blockIdx.x, blockIdx.y, gridDim.y, blockIdx.z, gridDim.z, gridDim.x * gridDim.y * gridDim.z);
}

// Fallback for frontiers whose per-tree histogram does not fit in shared memory: warp-aggregated
// global atomics.  Lanes that target the same counter are found with __match_any_sync; one leader
// per group issues a single atomic with the group's total.  grid = (chunks, T).
__global__ void __launch_bounds__(256) hist_global_kernel(const HistArgs a, int n_attrs,
                                                          const int32_t* bin_off,
                                                          const int32_t* wide) {
    const int t = blockIdx.y;
    const int nb = a.node_base[t];
    const uint16_t* nid = a.nid + (long long)t * a.Npad;
    const bool hw = a.wts != nullptr && a.has_w[t];
    const uint8_t* wt = hw ? a.wts + (long long)t * a.Npad : nullptr;
    const unsigned lane = threadIdx.x & 31;
    const long long stride = (long long)gridDim.x * blockDim.x;
    // loop bound rounded up to a whole warp so that every lane reaches __match_any_sync
    const long long nround = (a.N + 31) & ~31LL;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < nround; r += stride) {
        bool live = r < a.N;
        uint32_t node = live ? nid[r] : AVT_NODE_NONE;
        uint32_t w = live ? (hw ? wt[r] : 1u) : 0u;
        live = live && node != AVT_NODE_NONE && w != 0;
        if (live && a.skip) live = a.skip[nb + node] == 0;
        const uint32_t cl = live ? a.cls[r] : 0;
        if (a.bad && live && (node >= (uint32_t)(a.node_base[t + 1] - nb) || cl >= (uint32_t)a.C)) {
            *a.bad = 1;
            live = false;
        }
        for (int at = 0; at < n_attrs; ++at) {
            unsigned long long idx = ~0ULL - lane;  // unique dummy key for idle lanes
            if (live) {
                uint32_t code = wide[at] ? ((const uint16_t*)a.cols[at])[r] : ((const uint8_t*)a.cols[at])[r];
                if (a.bad && code >= (uint32_t)(bin_off[at + 1] - bin_off[at])) *a.bad = 1;
                else idx = ((unsigned long long)(nb + node) * a.TB + bin_off[at] + code) * a.C + cl;
            }
            const unsigned peers = __match_any_sync(0xffffffffu, idx);
            const uint32_t sum = __reduce_add_sync(peers, w);
            if (live && idx < ~0ULL - 31 && (unsigned)(__ffs(peers) - 1) == lane) atomicAdd(&a.H[idx], sum);
        }
    }
}

// One byte per row holding the codes of all (<= 4, narrow) attributes: attribute q at bit shift[q].
// One thread packs 16 rows; the shifted codes never leave their byte (the widths add up to <= 8 bits).
__global__ void pack_codes_kernel(const void* const* cols, int n_attrs, uint4 shifts, long long npad, uint8_t* out) {
    const long long base = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 16;
    if (base >= npad) return;
    const uint32_t sh[4] = {shifts.x, shifts.y, shifts.z, shifts.w};
    uint4 acc = make_uint4(0, 0, 0, 0);
    for (int q = 0; q < n_attrs; ++q) {
        const uint4 v = *reinterpret_cast<const uint4*>((const char*)cols[q] + base);
        acc.x |= v.x << sh[q]; acc.y |= v.y << sh[q]; acc.z |= v.z << sh[q]; acc.w |= v.w << sh[q];
    }
    *reinterpret_cast<uint4*>(out + base) = acc;
}

// ------------------------------------------------------------------------------------------------
// (b) impurity scan
// ------------------------------------------------------------------------------------------------
struct SplitTable {
    const avt_split* splits;    // [S]
    const int32_t* bounds;      // numeric upper bounds (inclusive bin index)
    const uint8_t* seg_lut;     // per-bin segment ids
    const int32_t* attr_first;  // [A+1] first split of each attribute (splits are grouped by attribute)
    const int32_t* bin_off;     // [A+1] first bin of each attribute in the full bin space
    int S, A, TB, C;
};

// Class counts of segment j of split sp.  `P` (shared, inclusive prefix over bins, [B][C]) is used
// when staged, otherwise the raw counts `Hn` ([B][C], global) of the attribute.
__device__ __forceinline__ void segment_counts(const SplitTable& st, const avt_split& sp, int j, int B,
                                               const uint32_t* P, const uint32_t* Hn, uint32_t* cnt) {
    const int C = st.C;
    for (int c = 0; c < C; ++c) cnt[c] = 0;
    if (sp.kind == AVT_SPLIT_NUMERIC) {
        const int lo = (j == 0) ? -1 : st.bounds[sp.off + j - 1];
        const int hi = (j == sp.n_seg - 1) ? B - 1 : st.bounds[sp.off + j];
        if (hi < 0 || hi <= lo) return;
        if (P) {
            for (int c = 0; c < C; ++c) cnt[c] = P[hi * C + c] - (lo >= 0 ? P[lo * C + c] : 0u);
        } else {
            for (int b = lo + 1; b <= hi; ++b)
                for (int c = 0; c < C; ++c) cnt[c] += Hn[b * C + c];
        }
    } else if (sp.kind == AVT_SPLIT_RANGES) {
        const int lo = st.bounds[sp.off + 2 * j], hi = min(st.bounds[sp.off + 2 * j + 1], B - 1);
        if (hi < lo) return;
        if (P) {
            for (int c = 0; c < C; ++c) cnt[c] = P[hi * C + c] - (lo > 0 ? P[(lo - 1) * C + c] : 0u);
        } else {
            for (int b = lo; b <= hi; ++b)
                for (int c = 0; c < C; ++c) cnt[c] += Hn[b * C + c];
        }
    } else {
        const uint8_t* lut = st.seg_lut + sp.off;
        for (int b = 0; b < B; ++b) {
            if (lut[b] != j) continue;
            if (P) {
                for (int c = 0; c < C; ++c) cnt[c] += P[b * C + c] - (b ? P[(b - 1) * C + c] : 0u);
            } else {
                for (int c = 0; c < C; ++c) cnt[c] += Hn[b * C + c];
            }
        }
    }
}

// Weighted-average impurity of one split at one node: BuilderReducer.expandTree's inner loop
// (tree/DecisionTreeBuilder.java:534-548).  Empty segments do not exist (App. B-10).  Terms are
// accumulated in segment order; for the two-segment splits in scope the sum is order-independent.
__device__ __forceinline__ double split_av(const SplitTable& st, const avt_split& sp, int B,
                                           const uint32_t* P, const uint32_t* Hn,
                                           const uint8_t* order, int criterion, avt_split_eval* ev) {
    uint32_t cnt[kMaxC];
    double weighted = 0.0;
    uint32_t total = 0;
    for (int j = 0; j < sp.n_seg; ++j) {
        segment_counts(st, sp, j, B, P, Hn, cnt);
        uint32_t n = 0;
        for (int c = 0; c < st.C; ++c) n += cnt[c];
        double stat = 0.0;
        if (n) {
            stat = impurity(cnt, st.C, order, criterion, n);
            weighted = __dadd_rn(weighted, __dmul_rn(stat, (double)n));
            total += n;
        }
        if (ev) {
            ev->seg_stat[j] = stat;
            for (int c = 0; c < st.C; ++c) ev->seg_count[j][c] = cnt[c];
        }
    }
    return __ddiv_rn(weighted, (double)total);
}

// Warp-cooperative inclusive prefix over the bins of one attribute: P[b][c] = sum_{b'<=b} H[b'][c].
__device__ __forceinline__ void warp_prefix(const uint32_t* Hn, uint32_t* P, int B, int C) {
    const unsigned lane = threadIdx.x & 31;
    for (int c = 0; c < C; ++c) {
        uint32_t carry = 0;
        for (int b0 = 0; b0 < B; b0 += 32) {
            const int b = b0 + lane;
            uint32_t v = b < B ? Hn[b * C + c] : 0u;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                uint32_t u = __shfl_up_sync(0xffffffffu, v, d);
                if (lane >= (unsigned)d) v += u;
            }
            if (b < B) P[b * C + c] = v + carry;
            carry += __shfl_sync(0xffffffffu, v, 31);
        }
    }
    __syncwarp();
}

struct ScanBest {
    double av;
    int split;
    int ties;
};

__device__ __forceinline__ void best_merge(ScanBest& a, const ScanBest& b) {
    if (b.split < 0) return;
    if (a.split < 0 || b.av < a.av) {
        a = b;
    } else if (b.av == a.av) {
        a.ties += b.ties;
        if (b.split < a.split) a.split = b.split;
    }
}

}  // namespace avt

#include "avt_scan.cuh"

namespace avt {

// Children statistics for chosen (node, split) pairs; one thread per pair (tiny).
__global__ void eval_kernel(const uint32_t* H, SplitTable st, const uint8_t* order, int criterion, int n,
                            const int32_t* nodes, const int32_t* splits, avt_split_eval* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const avt_split sp = st.splits[splits[i]];
    const int B = st.bin_off[sp.attr + 1] - st.bin_off[sp.attr];
    const uint32_t* Hn = H + ((long long)nodes[i] * st.TB + st.bin_off[sp.attr]) * st.C;
    avt_split_eval ev;
    for (int j = 0; j < kMaxSeg; ++j) {
        ev.seg_stat[j] = 0.0;
        for (int c = 0; c < kMaxC; ++c) ev.seg_count[j][c] = 0;
    }
    ev.av = split_av(st, sp, B, nullptr, Hn, order, criterion, &ev);
    out[i] = ev;
}

__global__ void class_counts_kernel(const uint32_t* H, int n_nodes, int TB, int C, int B0, uint32_t* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes * C) return;
    const int node = i / C, c = i % C;
    uint32_t s = 0;
    for (int b = 0; b < B0; ++b) s += H[((long long)node * TB + b) * C + c];
    out[i] = s;
}

// ClassPartitionGenerator statistics of every split for the rows of one node.
// util/AttributeSplitStat.java:153-170 (split info), 191-218 (gini/entropy), 240-282 (Hellinger),
// 298-336 + 452-471 (class-confidence ratio).  Segments iterate in ascending index (HashMap<Integer>
// order for < 16 segments); segments without rows do not exist.
// The statistics of one split from its segments' class counts (r.seg_count filled in): split information, then the criterion.
__device__ __noinline__ void cpg_statistics(const SplitTable& st, const avt_split& sp, const uint8_t* order, int criterion, avt_cpg_stat& r) {
    const int C = st.C;
    uint32_t segn[kMaxSeg];
    uint32_t total = 0;
    for (int j = 0; j < kMaxSeg; ++j) segn[j] = 0;
    for (int j = 0; j < sp.n_seg; ++j) {
        for (int c = 0; c < C; ++c) segn[j] += r.seg_count[j][c];
        total += segn[j];
    }
    const double log2v = 0.6931471805599453;
    const double tot = (double)total;
    // intrinsic split information
    double info = 0.0;
    for (int j = 0; j < sp.n_seg; ++j) {
        if (!segn[j]) continue;
        double pr = __ddiv_rn((double)segn[j], tot);
        info = __dsub_rn(info, __ddiv_rn(__dmul_rn(pr, strict_log(pr)), log2v));
    }
    r.split_info = info;
    double stat;
    if (criterion == AVT_CRIT_GINI || criterion == AVT_CRIT_ENTROPY) {
        double sum = 0.0;
        for (int j = 0; j < sp.n_seg; ++j) {
            if (!segn[j]) continue;
            double sj = impurity(r.seg_count[j], C, order, criterion, segn[j]);
            sum = __dadd_rn(sum, __dmul_rn(sj, (double)segn[j]));
        }
        stat = __ddiv_rn(sum, tot);
    } else if (criterion == AVT_CRIT_HELLINGER) {
        const int c0 = order[0], c1 = order[1];
        uint32_t n0 = 0, n1 = 0;
        for (int j = 0; j < sp.n_seg; ++j) {
            n0 += r.seg_count[j][c0];
            n1 += r.seg_count[j][c1];
        }
        double sum = 0.0;
        for (int j = 0; j < sp.n_seg; ++j) {
            if (!segn[j]) continue;
            double v0 = __dsqrt_rn(__ddiv_rn((double)r.seg_count[j][c0], (double)n0));
            double v1 = __dsqrt_rn(__ddiv_rn((double)r.seg_count[j][c1], (double)n1));
            double d = __dsub_rn(v0, v1);
            sum = __dadd_rn(sum, __dmul_rn(d, d));
        }
        stat = __dsqrt_rn(sum);
    } else {
        uint32_t ctot[kMaxC];
        for (int c = 0; c < C; ++c) {
            ctot[c] = 0;
            for (int j = 0; j < sp.n_seg; ++j) ctot[c] += r.seg_count[j][c];
        }
        double sum = 0.0;
        for (int j = 0; j < sp.n_seg; ++j) {
            if (!segn[j]) continue;
            double conf[kMaxC];
            double tc = 0.0;
            for (int i = 0; i < C; ++i) {
                const int c = order[i];
                conf[i] = __ddiv_rn((double)r.seg_count[j][c], (double)ctot[c]);
                tc = __dadd_rn(tc, conf[i]);
            }
            double ent = 0.0;
            for (int i = 0; i < C; ++i) {
                double ccr = __ddiv_rn(conf[i], tc);
                ent = __dsub_rn(ent, __ddiv_rn(__dmul_rn(ccr, strict_log(ccr)), log2v));
            }
            sum = __dadd_rn(sum, __dmul_rn(ent, (double)segn[j]));
        }
        stat = __ddiv_rn(sum, tot);
    }
    r.stat = stat;
}

// One WARP per split: the lanes stride over the attribute's bins to take the segments' class counts (coalesced reads of the
// per-bin segment table and of H; a split over a 1 000-level categorical is 1 000 bins x n_seg passes -- one thread per split
// spent 18 ms per sweep on C4's 81 860 splits), lane 0 then does the fp64 statistics in exactly the scalar order, and the
// 528-byte result goes out with coalesced stores.
constexpr int kCpgWarps = 8;
__global__ void __launch_bounds__(kCpgWarps * 32) cpg_kernel(const uint32_t* H, SplitTable st, int node, const uint8_t* order, int criterion,
                                                             avt_cpg_stat* out) {
    __shared__ avt_cpg_stat s_r[kCpgWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int s = blockIdx.x * kCpgWarps + warp;
    if (s >= st.S) return;   // whole warps leave; no block-wide barrier below
    const avt_split sp = st.splits[s];
    const int C = st.C;
    const int B = st.bin_off[sp.attr + 1] - st.bin_off[sp.attr];
    const uint32_t* Hn = H + ((long long)node * st.TB + st.bin_off[sp.attr]) * C;
    avt_cpg_stat& r = s_r[warp];
    {
        uint32_t* rc = &r.seg_count[0][0];
        for (int i = lane; i < kMaxSeg * kMaxC; i += 32) rc[i] = 0;
    }
    __syncwarp();
    for (int j = 0; j < sp.n_seg; ++j) {
        uint32_t cnt[kMaxC];
        for (int c = 0; c < kMaxC; ++c) cnt[c] = 0;
        int lo = 0, hi = -1;           // bins [lo, hi] of the segment (numeric / ranges); the table decides for the rest
        const uint8_t* lut = nullptr;
        if (sp.kind == AVT_SPLIT_NUMERIC) {
            lo = (j == 0) ? 0 : st.bounds[sp.off + j - 1] + 1;
            hi = (j == sp.n_seg - 1) ? B - 1 : st.bounds[sp.off + j];
            if (hi < 0 || hi < lo) { lo = 0; hi = -1; }
        } else if (sp.kind == AVT_SPLIT_RANGES) {
            lo = st.bounds[sp.off + 2 * j];
            hi = min(st.bounds[sp.off + 2 * j + 1], B - 1);
        } else {
            lut = st.seg_lut + sp.off;
            lo = 0;
            hi = B - 1;
        }
        for (int b = lo + lane; b <= hi; b += 32) {
            if (lut && lut[b] != j) continue;
            for (int c = 0; c < C; ++c) cnt[c] += Hn[b * C + c];
        }
        for (int c = 0; c < C; ++c) {
            uint32_t v = cnt[c];
#pragma unroll
            for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0) r.seg_count[j][c] = v;
        }
    }
    __syncwarp();
    if (lane == 0) cpg_statistics(st, sp, order, criterion, r);
    __syncwarp();
    const uint32_t* src = reinterpret_cast<const uint32_t*>(&r);
    uint32_t* dst = reinterpret_cast<uint32_t*>(out + s);
    for (int i = lane; i < (int)(sizeof(avt_cpg_stat) / 4); i += 32) dst[i] = src[i];
}


// ------------------------------------------------------------------------------------------------
// (c) repartition
// ------------------------------------------------------------------------------------------------
// nid[t][r] <- child_lut[(node_base[t]+nid[t][r]) * lut_width + code(attr of that node, r)].
// One thread per 8 rows (one 128-bit load/store of node ids); the gathered code bytes of a warp
// fall in at most (#distinct split attributes among its rows) 256-byte spans.
__global__ void __launch_bounds__(256) repartition_kernel(uint16_t* nid_all, const void* const* cols,
                                                          const int32_t* wide, const int32_t* node_base,
                                                          const int32_t* node_attr, const uint16_t* child_lut,
                                                          int lut_width, long long N, long long Npad) {
    const int t = blockIdx.y;
    uint16_t* nid = nid_all + (long long)t * Npad;
    const int nb = node_base[t];
    const long long stride = (long long)gridDim.x * blockDim.x * 8;
    for (long long base = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8; base < N; base += stride) {
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
                    nw = child_lut[(long long)g * lut_width + code];
                }
            }
            wds[j >> 1] = (wds[j >> 1] & ~(0xFFFFu << ((j & 1) * 16))) | (nw << ((j & 1) * 16));
        }
        *reinterpret_cast<uint4*>(nid + base) = make_uint4(wds[0], wds[1], wds[2], wds[3]);
    }
}

// The same for narrow schemas (<= 4 uint8 attributes) and frontiers whose tables fit shared memory: the tables are
// staged once per block and a thread moves 16 rows with vector loads and the byte-permute selection of
// repart16_fast -- no per-row gathers.  grid = (chunks, T), dynamic smem = (n_max + 2) * (4 + 2 * lut_w) + 8.
struct RepartFastArgs {
    uint16_t* nid;             // [T][Npad], updated in place
    const void* const* cols;   // [A]
    const uint8_t* packed;     // bit-packed codes (HistArgs::packed) or nullptr
    uint32_t pk_shift[4], pk_mask4[4];
    const int32_t* node_base;  // [T+1] frontier the ids refer to
    const int32_t* node_attr;
    const uint16_t* child_lut;
    int lut_w, n_attrs;
    long long N, Npad;
};

__global__ void __launch_bounds__(256) repartition_fast_kernel(const RepartFastArgs a) {
    extern __shared__ uint32_t rsh[];
    const int t = blockIdx.y;
    const int nb = a.node_base[t], n_prev = a.node_base[t + 1] - nb;
    uint32_t* sel = rsh;
    uint16_t* lut = reinterpret_cast<uint16_t*>(sel + n_prev + 1);
    for (int i = threadIdx.x; i <= n_prev; i += blockDim.x) {
        const int at = i < n_prev ? a.node_attr[nb + i] : -1;
        sel[i] = 0x4440u | (uint32_t)(at < 0 ? 0 : at);
    }
    for (int i = threadIdx.x; i < (n_prev + 1) * a.lut_w; i += blockDim.x)
        lut[i] = i < n_prev * a.lut_w ? a.child_lut[(long long)nb * a.lut_w + i] : (uint16_t)AVT_NODE_NONE;
    __syncthreads();
    const RepartSmem rp{sel, lut, (uint32_t)n_prev, true, nullptr};
    uint16_t* nid = a.nid + (long long)t * a.Npad;
    const long long stride = (long long)gridDim.x * blockDim.x * kRowsPerThread;
    for (long long base = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * kRowsPerThread; base < a.N; base += stride) {
        const uint4 vn0 = ld_stream(nid + base), vn1 = ld_stream(nid + base + 8);
        uint4 v[4];
        if (a.packed) {
            const uint4 vp = ld_stream(a.packed + base);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const uint32_t sh = a.pk_shift[q], m = q < a.n_attrs ? a.pk_mask4[q] : 0u;
                v[q] = make_uint4((vp.x >> sh) & m, (vp.y >> sh) & m, (vp.z >> sh) & m, (vp.w >> sh) & m);
            }
        } else {
#pragma unroll
            for (int q = 0; q < 4; ++q) v[q] = q < a.n_attrs ? ld_stream((const char*)a.cols[q] + base) : make_uint4(0, 0, 0, 0);
        }
        uint32_t nw[8];
        repart16_fast(vn0, vn1, v, 4, rp, a.lut_w, nw);
        *reinterpret_cast<uint4*>(nid + base) = make_uint4(nw[0], nw[1], nw[2], nw[3]);
        *reinterpret_cast<uint4*>(nid + base + 8) = make_uint4(nw[4], nw[5], nw[6], nw[7]);
    }
}

}  // namespace avt

#include "avt_scan_fast.cuh"
#include "avt_forest.cuh"
#include "avt_joint.cuh"
#include "avt_sorted.cuh"
#include "avt_csv.cuh"
#include "avt_predict.cuh"
#include "avt_bagging.cuh"
