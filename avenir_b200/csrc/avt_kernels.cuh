// Hand-written sm_100a kernels of the split-evaluation path.
//   (a) hist_smem_kernel / hist_global_kernel : (tree,node,attr,bin,class) count histograms
//   (b) scan_kernel / eval_kernel / cpg_kernel: fused fp64 impurity scan + per-node arg-min
//   (c) repartition_kernel                    : row -> child node
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
    const uint16_t* nid;       // [T][Npad]
    const uint8_t* wts;        // [T][Npad] or nullptr
    const uint8_t* has_w;      // [T] (nullptr when wts == nullptr)
    uint32_t* H;               // [NF][TB][C]
    const int32_t* node_base;  // [T+1]
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
};

template <bool HAS_W>
__device__ __forceinline__ void hist_rows16(uint32_t* sh, const HistArgs& a, int e0, int e1,
                                            long long base, const uint16_t* nid, const uint8_t* wt,
                                            int GB, int rlog, uint32_t rep) {
    const uint4 vn0 = ld_stream(nid + base);
    const uint4 vn1 = ld_stream(nid + base + 8);
    const uint4 vc = ld_stream(a.cls + base);
    uint4 vw = make_uint4(0x01010101u, 0x01010101u, 0x01010101u, 0x01010101u);
    if (HAS_W) vw = ld_stream(wt + base);
    // per-row base key (node*GB)*C + cls ; rows outside the frontier get a negative marker
    int rowkey[kRowsPerThread];
    uint32_t roww[kRowsPerThread];
#pragma unroll
    for (int j = 0; j < kRowsPerThread; ++j) {
        uint32_t node = j < 8 ? u4_half(vn0, j) : u4_half(vn1, j - 8);
        uint32_t w = u4_byte(vw, j);
        roww[j] = w;
        rowkey[j] = (node == AVT_NODE_NONE || w == 0) ? -1 : (int)(node * GB * a.C + u4_byte(vc, j));
    }
    for (int e = e0; e < e1; e += 2) {
        // two columns in flight per iteration
        const bool two = (e + 1 < e1);
        const int wide0 = a.ent_wide[e], wide1 = two ? a.ent_wide[e + 1] : 0;
        const char* c0 = (const char*)a.cols[a.ent_attr[e]];
        const char* c1 = two ? (const char*)a.cols[a.ent_attr[e + 1]] : c0;
        uint4 v0a, v0b, v1a, v1b;
        if (wide0) { v0a = ld_stream(c0 + base * 2); v0b = ld_stream(c0 + base * 2 + 16); }
        else { v0a = ld_stream(c0 + base); v0b = v0a; }
        if (two) {
            if (wide1) { v1a = ld_stream(c1 + base * 2); v1b = ld_stream(c1 + base * 2 + 16); }
            else { v1a = ld_stream(c1 + base); v1b = v1a; }
        } else { v1a = v0a; v1b = v0b; }
        const int off0 = a.ent_locoff[e] * a.C;
        const int off1 = two ? a.ent_locoff[e + 1] * a.C : 0;
#pragma unroll
        for (int j = 0; j < kRowsPerThread; ++j) {
            if (rowkey[j] < 0) continue;
            uint32_t code0 = wide0 ? (j < 8 ? u4_half(v0a, j) : u4_half(v0b, j - 8)) : u4_byte(v0a, j);
            uint32_t k0 = (uint32_t)(rowkey[j] + off0 + code0 * a.C);
            atomicAdd(&sh[(k0 << rlog) + rep], roww[j]);
            if (two) {
                uint32_t code1 = wide1 ? (j < 8 ? u4_half(v1a, j) : u4_half(v1b, j - 8)) : u4_byte(v1a, j);
                uint32_t k1 = (uint32_t)(rowkey[j] + off1 + code1 * a.C);
                atomicAdd(&sh[(k1 << rlog) + rep], roww[j]);
            }
        }
    }
}

// Shared-memory privatised histogram.  grid = (T, chunks, G).  Block (t, chunk, g) counts the rows of
// its grid-strided 16-row vectors into a private copy of H[t][*][bins of group g][*] kept in shared
// memory, replicated 2^rlog times (copy = lane mod replicas) so that lanes of a warp that hit the
// same counter land in different banks; the block then folds the copies and adds them to global H.
template <bool HAS_W>
__global__ void __launch_bounds__(kHistThreads) hist_smem_kernel(const HistArgs a) {
    extern __shared__ uint32_t sh[];
    const int t = blockIdx.x, g = blockIdx.z;
    const int nb = a.node_base[t], n_nodes = a.node_base[t + 1] - nb;
    const int GB = a.grp_bins[g], rlog = a.grp_rlog[g];
    const int keys = n_nodes * GB * a.C;
    const int words = keys << rlog;
    for (int i = threadIdx.x; i < words; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    const uint16_t* nid = a.nid + (long long)t * a.Npad;
    const bool tree_has_w = HAS_W && a.has_w[t];
    const uint8_t* wt = HAS_W ? a.wts + (long long)t * a.Npad : nullptr;
    const uint32_t rep = (threadIdx.x & 31) & ((1u << rlog) - 1);
    const int e0 = a.grp_start[g], e1 = a.grp_start[g + 1];
    const long long stride = (long long)gridDim.y * blockDim.x * kRowsPerThread;
    for (long long base = ((long long)blockIdx.y * blockDim.x + threadIdx.x) * kRowsPerThread;
         base < a.N; base += stride) {
        if (tree_has_w) hist_rows16<true>(sh, a, e0, e1, base, nid, wt, GB, rlog, rep);
        else hist_rows16<false>(sh, a, e0, e1, base, nid, wt, GB, rlog, rep);
    }
    __syncthreads();
    const int* l2g = a.grp_l2g + a.grp_l2g_off[g];
    const int R = 1 << rlog;
    for (int k = threadIdx.x; k < keys; k += blockDim.x) {
        uint32_t s = 0;
        for (int r = 0; r < R; ++r) s += sh[(k << rlog) + ((r + threadIdx.x) & (R - 1))];
        if (s) {
            const int c = k % a.C;
            const int lb = (k / a.C) % GB;
            const int node = k / (a.C * GB);
            atomicAdd(&a.H[((long long)(nb + node) * a.TB + l2g[lb]) * a.C + c], s);
        }
    }
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
        const uint32_t cl = live ? a.cls[r] : 0;
        for (int at = 0; at < n_attrs; ++at) {
            unsigned long long idx = ~0ULL - lane;  // unique dummy key for idle lanes
            if (live) {
                uint32_t code = wide[at] ? ((const uint16_t*)a.cols[at])[r] : ((const uint8_t*)a.cols[at])[r];
                idx = ((unsigned long long)(nb + node) * a.TB + bin_off[at] + code) * a.C + cl;
            }
            const unsigned peers = __match_any_sync(0xffffffffu, idx);
            const uint32_t sum = __reduce_add_sync(peers, w);
            if (live && (unsigned)(__ffs(peers) - 1) == lane) atomicAdd(&a.H[idx], sum);
        }
    }
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

// One block per frontier node.  Warps take the node's candidate attributes round-robin; a warp
// stages the attribute's prefix counts in shared memory, its lanes evaluate the attribute's splits
// and write av[node][split]; the block reduces to (min av, lowest split index, number of ties).
__global__ void __launch_bounds__(kScanThreads) scan_kernel(const uint32_t* H, SplitTable st,
                                                            const uint32_t* cand_mask,
                                                            const uint8_t* order, int criterion,
                                                            double* av_buf, avt_node_best* out) {
    __shared__ uint32_t Pbuf[kScanWarps][kScanWordsPerWarp];
    __shared__ ScanBest wbest[kScanWarps];
    const int node = blockIdx.x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int C = st.C;
    const uint32_t* Hnode = H + (long long)node * st.TB * C;
    const int mask_words = (st.A + 31) >> 5;
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    ScanBest best{0.0, -1, 0};
    for (int at = warp; at < st.A; at += kScanWarps) {
        const int s0 = st.attr_first[at], s1 = st.attr_first[at + 1];
        if (s0 == s1) continue;
        const bool cand = cand_mask == nullptr ||
                          ((cand_mask[(long long)node * mask_words + (at >> 5)] >> (at & 31)) & 1u);
        if (!cand) {
            for (int s = s0 + lane; s < s1; s += 32) av_buf[(long long)node * st.S + s] = nan;
            continue;
        }
        const int B = st.bin_off[at + 1] - st.bin_off[at];
        const uint32_t* Hn = Hnode + (long long)st.bin_off[at] * C;
        const uint32_t* P = nullptr;
        if (B * C <= kScanWordsPerWarp) {
            __syncwarp();
            warp_prefix(Hn, Pbuf[warp], B, C);
            P = Pbuf[warp];
        }
        for (int s = s0 + lane; s < s1; s += 32) {
            const avt_split sp = st.splits[s];
            const double av = split_av(st, sp, B, P, Hn, order, criterion, nullptr);
            av_buf[(long long)node * st.S + s] = av;
            ScanBest b{av, s, 1};
            if (av == av) best_merge(best, b);  // NaN never wins (`av < min` is false in Java too)
        }
    }
    // warp reduce
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        ScanBest o;
        o.av = __shfl_xor_sync(0xffffffffu, best.av, d);
        o.split = __shfl_xor_sync(0xffffffffu, best.split, d);
        o.ties = __shfl_xor_sync(0xffffffffu, best.ties, d);
        best_merge(best, o);
    }
    if (lane == 0) wbest[warp] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        ScanBest b = wbest[0];
        for (int w = 1; w < kScanWarps; ++w) best_merge(b, wbest[w]);
        // the node's own class counts: bins of the first attribute that has any
        uint32_t cnt[kMaxC];
        for (int c = 0; c < C; ++c) cnt[c] = 0;
        const int B0 = st.bin_off[1] - st.bin_off[0];
        for (int bb = 0; bb < B0; ++bb)
            for (int c = 0; c < C; ++c) cnt[c] += Hnode[bb * C + c];
        uint32_t n = 0;
        for (int c = 0; c < C; ++c) n += cnt[c];
        avt_node_best r;
        r.min_av = b.split >= 0 ? b.av : nan;
        r.node_stat = impurity(cnt, C, order, criterion, n);
        r.population = n;
        r.best_split = b.split;
        r.n_ties = b.ties;
        out[node] = r;
    }
}

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
__global__ void cpg_kernel(const uint32_t* H, SplitTable st, int node, const uint8_t* order, int criterion,
                           avt_cpg_stat* out) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= st.S) return;
    const avt_split sp = st.splits[s];
    const int C = st.C;
    const int B = st.bin_off[sp.attr + 1] - st.bin_off[sp.attr];
    const uint32_t* Hn = H + ((long long)node * st.TB + st.bin_off[sp.attr]) * C;
    avt_cpg_stat r;
    uint32_t segn[kMaxSeg];
    uint32_t total = 0;
    for (int j = 0; j < kMaxSeg; ++j) {
        segn[j] = 0;
        for (int c = 0; c < kMaxC; ++c) r.seg_count[j][c] = 0;
    }
    for (int j = 0; j < sp.n_seg; ++j) {
        uint32_t cnt[kMaxC];
        segment_counts(st, sp, j, B, nullptr, Hn, cnt);
        for (int c = 0; c < C; ++c) {
            r.seg_count[j][c] = cnt[c];
            segn[j] += cnt[c];
        }
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
    out[s] = r;
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

}  // namespace avt
