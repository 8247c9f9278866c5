// Kernel (b), latency-optimised form for small split tables (the level loop of narrow schemas such as
// call_hangup.json is bound by the serial dependency chain of this kernel, not by its throughput):
//   * everything that depends only on the split table -- the thread's (split, segment) item, its bounds or
//     bin mask, the class order -- is loaded BEFORE the kernel waits for the histogram pass (programmatic
//     dependent launch), so it costs nothing on the critical path;
//   * one thread per (split, segment) evaluates that segment's impurity, instead of one thread per split
//     walking its segments one after the other; one thread per split then combines the segments in order
//     with exactly the operations of split_av, so the results are bit-identical;
//   * the children of the best split are read back from those per-segment results instead of being recomputed;
//   * the part of the level planning that concerns one node only (split attribute, number of children, the
//     bin -> child row with child ranks LOCAL to the node, the node's record) is done by the node's own block;
//     the block that finishes last only turns the local ranks into frontier ids with a prefix sum.
// Same inputs and outputs as scan_kernel + plan_level (node_attr, child_lut, next_count, next_node_base,
// records); av / best / best_eval are not produced (nothing in the level loop reads them).
// Included by avt_kernels.cuh after avt_scan.cuh.
#pragma once

namespace avt {

constexpr int kFastItems = kScanThreads;   // (split, segment) pairs: one per thread
constexpr int kFastCntWords = 2048;        // class counts of all items

// Extra device tables of the split table, built by avt_split_table_set.
struct FastTable {
    const uint32_t* items;     // [n_items] (split << 4) | segment, splits in table order, segments ascending
    const int32_t* seg_first;  // [S+1] first item of every split
    int32_t* nkids;            // [NF] children of every node of the current frontier (scratch between the blocks and the planner)
    int n_items;
    // look-back planning (frontier bound <= kLookbackNodes): every node's block publishes its child count tagged with this
    // launch's epoch and reads the counts of the nodes before it, so no block has to wait for the whole grid
    const uint32_t* cand;      // [NF] candidate attributes of every node (one word: the table has <= 8 attributes), nullptr = all
    uint32_t* pub;             // [kLookbackNodes], nullptr = the last block plans the level (ticket)
    uint32_t epoch;            // 1 .. 2^24-1, different from the tags any earlier launch left in `pub`
};

constexpr int kLookbackNodes = 512;
constexpr int kPushWords = 1024;    // push form of the peer exchange: counters per node (bins x classes) at most
constexpr int kPushCap = 1 << 16;   // ... and counters per level (receive slots per source rank and buffer)

__device__ __forceinline__ void st_relaxed_gpu(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ uint16_t ldcg_u16(const uint16_t* p) { return __ldcg(p); }

// The planning that is left for the last block: per tree, an exclusive prefix sum over the nodes' child counts
// turns node-local child ranks into ids of the next frontier.
__device__ void plan_level_fast(const PlanArgs& p, const int32_t* nkids, int* warp_tot, int* running, int* s_first) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int t = 0; t < p.n_trees; ++t) {
        const int nb = p.node_base[t], n = p.node_base[t + 1] - nb;
        if (threadIdx.x == 0) *running = 0;
        __syncthreads();
        for (int i0 = 0; i0 < n; i0 += blockDim.x) {
            const int i = i0 + threadIdx.x;
            const int nk = i < n ? ldcg_i32(nkids + nb + i) : 0;
            int v = nk;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int u = __shfl_up_sync(0xffffffffu, v, d);
                if (lane >= d) v += u;
            }
            if (lane == 31) warp_tot[warp] = v;
            __syncthreads();
            int woff = 0, tot = 0;
            for (int w = 0; w < kScanWarps; ++w) {
                if (w < warp) woff += warp_tot[w];
                tot += warp_tot[w];
            }
            const int first = *running + woff + v - nk;
            s_first[threadIdx.x] = first;
            if (nk && first + nk > (int)AVT_NODE_NONE - 1) *p.err = 1;
            __syncthreads();
            const int cnt = min((int)blockDim.x, n - i0);
            uint16_t* lut = p.child_lut + (long long)(nb + i0) * p.lut_w;
            for (int e = threadIdx.x; e < cnt * p.lut_w; e += blockDim.x) {
                const uint16_t c = ldcg_u16(lut + e);
                if (c != (uint16_t)AVT_NODE_NONE) lut[e] = (uint16_t)(c + s_first[e / p.lut_w]);
            }
            if (p.records) {
                for (int e = threadIdx.x; e < cnt * kMaxSeg; e += blockDim.x) {
                    int32_t* c = &p.records[nb + i0 + e / kMaxSeg].child[e % kMaxSeg];
                    const int32_t r = ldcg_i32(c);
                    if (r >= 0) *c = r + s_first[e / kMaxSeg];
                }
            }
            __syncthreads();
            if (threadIdx.x == 0) *running += tot;
            __syncthreads();
        }
        if (threadIdx.x == 0) p.next_count[t] = *running;
        __syncthreads();
    }
    if (threadIdx.x == 0 && p.next_node_base) {
        int acc = 0;
        for (int k = 0; k < p.n_trees; ++k) {
            p.next_node_base[k] = acc;
            acc += p.next_count[k];
        }
        p.next_node_base[p.n_trees] = acc;
    }
    if (p.zero_ptr)
        for (long long i = threadIdx.x; i < p.zero_words; i += blockDim.x) p.zero_ptr[i] = 0;
}

constexpr int kFastWordsPerAttr = 512;     // prefix counts of one attribute ([bins][classes])

// grid = frontier bound (blocks beyond the true frontier only take part in the last-block election),
// block = kScanThreads.  Requires (checked by the host, see fast_scan_ok): A <= kScanWarps, S <= kScanThreads,
// n_items <= kFastItems, n_items * C <= kFastCntWords, bins * C <= kFastWordsPerAttr for every attribute with
// splits, <= 64 bins for attributes with segment-table splits.
__global__ void __launch_bounds__(kScanThreads) scan_fast_kernel(const uint32_t* H, SplitTable st, FastTable ft, const uint8_t* order_g,
                                                                 int criterion, const int32_t* n_nodes_dev, PlanArgs plan) {
    __shared__ uint32_t Pbuf[kScanWarps][kFastWordsPerAttr];
    __shared__ double it_stat[kFastItems];
    __shared__ uint32_t it_n[kFastItems];
    __shared__ uint32_t it_cnt[kFastCntWords];
    __shared__ avt_split s_splits[kScanThreads];
    __shared__ int32_t s_seg_first[kScanThreads + 1];
    __shared__ ScanBest wbest[kScanWarps];
    __shared__ uint8_t order[kMaxC];
    __shared__ uint32_t s_cnt[kMaxC];
    __shared__ uint32_t s_pop;
    __shared__ avt_grow_record s_rec;
    __shared__ int s_warp_tot[kScanWarps];
    __shared__ int s_running;
    __shared__ int s_is_last;
    __shared__ int s_first[kScanThreads];
    __shared__ uint32_t s_sum[kPushWords];   // push exchange: the node's counts summed over the ranks
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int C = st.C;
    pdl_launch_dependents();
    // ---- before the wait: the thread's item and split, the class order, the warp's attribute
    int my_attr = -1, my_kind = 0, my_lo = -1, my_hi = -1;
    unsigned long long my_mask = 0;
    if (tid < ft.n_items) {
        const uint32_t it = ft.items[tid];
        const avt_split sp = st.splits[it >> 4];
        const int j = (int)(it & 15u);
        const int B = st.bin_off[sp.attr + 1] - st.bin_off[sp.attr];
        my_attr = sp.attr;
        my_kind = sp.kind;
        if (sp.kind == AVT_SPLIT_NUMERIC) {
            my_lo = (j == 0) ? -1 : st.bounds[sp.off + j - 1];
            my_hi = (j == sp.n_seg - 1) ? B - 1 : st.bounds[sp.off + j];
        } else {
            const uint8_t* lut = st.seg_lut + sp.off;
            for (int b = 0; b < B; ++b)
                if (lut[b] == j) my_mask |= 1ull << b;
        }
    }
    if (tid < st.S) s_splits[tid] = st.splits[tid];
    if (tid <= st.S) s_seg_first[tid] = ft.seg_first[tid];
    if (tid == kScanThreads - 1 && st.S == kScanThreads) s_seg_first[st.S] = ft.seg_first[st.S];
    if (tid < C) order[tid] = order_g[tid];
    int w_B = 0, w_off = 0;
    bool w_has = false;
    if (warp < st.A) {
        w_off = st.bin_off[warp];
        w_B = st.bin_off[warp + 1] - w_off;
        w_has = st.attr_first[warp + 1] > st.attr_first[warp];
    }
    const int B0 = st.bin_off[1] - st.bin_off[0];
    pdl_wait();   // the histogram (and the frontier) come from the previous launch
    if (plan.signal_flags && blockIdx.x == 0 && tid < plan.n_ranks)
        st_release_sys(plan.signal_flags[tid], plan.peer_epoch);   // the pass is complete: its counts are in this GPU's L2
    __syncthreads();
    const int node = blockIdx.x;
    const bool lookback = ft.pub != nullptr;
    if (plan.peer_H) {
        // wait until every rank has finished this level's histogram pass (which also means: its previous scan)
        if (tid < plan.n_ranks)
            if (!wait_flag_bounded(plan.peer_flags + tid, plan.peer_epoch)) *plan.err = 3;
        __syncthreads();
    }
    const int n_nodes = *n_nodes_dev;
    const bool push = plan.push_dst != nullptr;
    if (push && n_nodes > 0) {
        // Push form of the exchange.  Tags are the level's epoch, so the slots never need clearing; slot buffers alternate
        // by level like the histogram buffers.  A rank's words for a node arrive only after that rank finished its previous
        // scan, so "all ranks' words are here" also licenses clearing the next level's histogram buffer below.
        const int words = st.TB * C;
        const int src_node = node < n_nodes ? node : 0;             // blocks without a node only wait (on node 0's first word)
        const long long off = (long long)src_node * words;
        const int others = plan.n_ranks - 1;
        const unsigned long long tag = (unsigned long long)plan.peer_epoch << 32;
        if (node < n_nodes) {
            for (int q = tid; q < words * others; q += blockDim.x) {
                const int w = q / others, k = q - w * others;
                const int d = k + (k >= plan.my_rank);
                st_relaxed_sys_u64(plan.push_dst[d] + off + w, tag | __ldcg(plan.H_local + off + w));
            }
            for (int w = tid; w < words; w += blockDim.x) s_sum[w] = __ldcg(plan.H_local + off + w);
            __syncthreads();
        }
        const int n_poll = node < n_nodes ? words * others : others;
        bool timeout = false;
        for (int q = tid; q < n_poll; q += blockDim.x) {
            const int w = node < n_nodes ? q / others : 0, k = node < n_nodes ? q - w * others : q;
            const int src = k + (k >= plan.my_rank);
            const unsigned long long* slot = plan.push_recv + (long long)src * plan.push_cap + off + w;
            unsigned long long v = 0;
            for (int spin = 0;; ++spin) {
                v = ld_relaxed_sys_u64(slot);
                if ((v >> 32) == plan.peer_epoch) break;
                if (spin > (1 << 24)) { timeout = true; break; }   // never hang the GPU: report instead
            }
            if (node < n_nodes) atomicAdd(&s_sum[w], (uint32_t)v);
        }
        if (timeout) *plan.err = 3;
        __syncthreads();
        if (node < n_nodes) {
            for (int w = tid; w < words; w += blockDim.x) plan.h_sum[off + w] = s_sum[w];
            __syncthreads();
        }
    }
    if (lookback && plan.zero_ptr) {
        // every block clears its share of the next level's histogram region (nobody reads that buffer any more:
        // it held the counts of the level before this one)
        const long long per = (plan.zero_words + gridDim.x - 1) / gridDim.x;
        const long long z1 = min(plan.zero_words, (long long)(blockIdx.x + 1) * per);
        for (long long i = (long long)blockIdx.x * per + tid; i < z1; i += blockDim.x) plan.zero_ptr[i] = 0;
    }
    if (node < n_nodes) {
        if (plan.peer_H) {
            // every rank's partial counts for this node, pulled over NVLink and summed (see scan_node)
            const long long words = (long long)st.TB * C, off = (long long)node * words;
            for (long long i = tid; i < words; i += blockDim.x) {
                uint32_t sum = 0;
                for (int r0 = 0; r0 < plan.n_ranks; r0 += 8) {   // 8 NVLink loads in flight per thread
                    uint32_t v[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) v[k] = r0 + k < plan.n_ranks ? ld_relaxed_sys(plan.peer_H[r0 + k] + off + i) : 0u;
#pragma unroll
                    for (int k = 0; k < 8; ++k) sum += v[k];
                }
                plan.h_sum[off + i] = sum;
            }
            __syncthreads();
        }
        const uint32_t* Hnode = H + (long long)node * st.TB * C;
        const double nan = __longlong_as_double(0x7ff8000000000000LL);
        // ---- phase 1: warp a stages the prefix counts of attribute a; the last warp adds the node's row count
        if (w_has) warp_prefix(Hnode + (long long)w_off * C, Pbuf[warp], w_B, C);
        if (warp == kScanWarps - 1) {
            if (lane < C) {
                uint32_t s = 0;
                for (int bb = 0; bb < B0; ++bb) s += Hnode[bb * C + lane];
                s_cnt[lane] = s;
            }
            __syncwarp();
            if (lane == 0) {
                uint32_t n = 0;
                for (int c = 0; c < C; ++c) n += s_cnt[c];
                s_pop = n;
            }
        }
        __syncthreads();
        // ---- phase 2: one thread per (split, segment): class counts from the prefix table, impurity
        if (tid < ft.n_items) {
            const uint32_t* P = Pbuf[my_attr];
            uint32_t* cnt = it_cnt + tid * C;
            uint32_t n = 0;
            if (my_kind == AVT_SPLIT_NUMERIC) {
                const bool some = !(my_hi < 0 || my_hi <= my_lo);
                for (int c = 0; c < C; ++c) {
                    const uint32_t v = some ? P[my_hi * C + c] - (my_lo >= 0 ? P[my_lo * C + c] : 0u) : 0u;
                    cnt[c] = v;
                    n += v;
                }
            } else {
                for (int c = 0; c < C; ++c) {
                    uint32_t v = 0;
                    for (unsigned long long m = my_mask; m; m &= m - 1) {
                        const int b = __ffsll((long long)m) - 1;
                        v += P[b * C + c] - (b ? P[(b - 1) * C + c] : 0u);
                    }
                    cnt[c] = v;
                    n += v;
                }
            }
            it_n[tid] = n;
            it_stat[tid] = n ? impurity(cnt, C, order, criterion, n) : 0.0;
        }
        __syncthreads();
        // ---- phase 3: one thread per split combines its segments in order (the operations of split_av)
        ScanBest best{0.0, -1, 0};
        if (tid < st.S) {
            const int i0 = s_seg_first[tid], i1 = s_seg_first[tid + 1];
            double weighted = 0.0;
            uint32_t total = 0;
            for (int i = i0; i < i1; ++i) {
                const uint32_t n = it_n[i];
                if (n) {
                    weighted = __dadd_rn(weighted, __dmul_rn(it_stat[i], (double)n));
                    total += n;
                }
            }
            const double av = __ddiv_rn(weighted, (double)total);
            const bool cand = ft.cand == nullptr || ((ft.cand[node] >> s_splits[tid].attr) & 1u);
            if (av == av && cand) best = ScanBest{av, tid, 1};   // NaN never wins
        }
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
        // ---- phase 4 (warp 0): block arg-min, then this node's share of the level planning
        if (warp == 0) {
            ScanBest b = wbest[0];
            for (int w = 1; w < kScanWarps; ++w) best_merge(b, wbest[w]);
            const uint32_t n = s_pop;
            const bool live = n > 0 && b.split >= 0;
            avt_split sp{};
            int i0 = 0;
            if (live) {
                sp = s_splits[b.split];
                i0 = s_seg_first[b.split];
            }
            // children: segment `lane` of the best split; a child exists where the segment has rows; its rank
            // among the node's children is the number of non-empty segments before it
            const bool seg = live && lane < sp.n_seg;
            const uint32_t pop = seg ? it_n[i0 + lane] : 0u;
            const unsigned has_mask = __ballot_sync(0xffffffffu, pop != 0);
            const int rank = __popc(has_mask & ((1u << lane) - 1u));
            const int nk = __popc(has_mask);
            // tree of this node
            int t = 0;
            for (int t0 = 0; t0 < plan.n_trees; t0 += 32) {
                const int tt = t0 + lane;
                const bool mine = tt < plan.n_trees && node >= plan.node_base[tt] && node < plan.node_base[tt + 1];
                const unsigned m = __ballot_sync(0xffffffffu, mine);
                if (m) { t = t0 + __ffs(m) - 1; break; }
            }
            int first = 0;   // id of this node's first child in the next frontier (look-back mode; else added by the planner)
            if (lookback) {
                if (lane == 0) st_relaxed_gpu(ft.pub + node, (ft.epoch << 8) | (uint32_t)nk);
                // child counts of all nodes before this one: before its tree / inside its tree
                const int nbt = plan.node_base[t];
                int before = 0, inside = 0;
                bool timeout = false;
                for (int i = lane; i < node; i += 32) {
                    uint32_t v = 0;
                    int spin = 0;
                    for (;;) {
                        v = ld_relaxed_gpu(ft.pub + i);
                        if ((v >> 8) == ft.epoch) break;
                        if (++spin > (1 << 22)) { timeout = true; break; }   // never hang the GPU: report instead
                    }
                    if (i < nbt) before += (int)(v & 0xFFu);
                    else inside += (int)(v & 0xFFu);
                }
#pragma unroll
                for (int d = 16; d; d >>= 1) {
                    before += __shfl_xor_sync(0xffffffffu, before, d);
                    inside += __shfl_xor_sync(0xffffffffu, inside, d);
                }
                first = inside;
                if (__any_sync(0xffffffffu, timeout) && lane == 0) *plan.err = 2;
                if (lane == 0 && nk && first + nk > (int)AVT_NODE_NONE - 1) *plan.err = 1;
                // next level's node_base: entry k is the number of children of all nodes before node_base[k]
                const int total = before + inside + nk;
                for (int k = lane; k <= plan.n_trees; k += 32) {
                    const int nbk = plan.node_base[k];
                    if (nbk == node + 1) plan.next_node_base[k] = total;
                    if (nbk == 0 && node == 0) plan.next_node_base[k] = 0;
                }
            }
            // bin -> child row (node-local ranks + first)
            const int Bs = live ? st.bin_off[sp.attr + 1] - st.bin_off[sp.attr] : 0;
            uint16_t* lut = plan.child_lut + (long long)node * plan.lut_w;
            for (int b0 = 0; b0 < plan.lut_w; b0 += 32) {
                const int bin = b0 + lane;
                int sg = -1;
                if (bin < Bs) {
                    if (sp.kind == AVT_SPLIT_NUMERIC) {
                        sg = 0;
                        while (sg < sp.n_seg - 1 && bin > st.bounds[sp.off + sg]) ++sg;
                    } else {
                        sg = st.seg_lut[sp.off + bin];
                        if (sg >= sp.n_seg) sg = -1;
                    }
                }
                const int src = sg < 0 ? 0 : sg;
                const int r = __shfl_sync(0xffffffffu, rank, src);
                const bool ok = sg >= 0 && ((has_mask >> src) & 1u);
                if (bin < plan.lut_w) lut[bin] = ok ? (uint16_t)(r + first) : (uint16_t)AVT_NODE_NONE;
            }
            if (lane == 0) {
                plan.node_attr[node] = live ? sp.attr : -1;
                ft.nkids[node] = nk;
            }
            if (plan.records) {
                for (int i = lane; i < (int)(sizeof(avt_grow_record) / 4); i += 32) reinterpret_cast<uint32_t*>(&s_rec)[i] = 0;
                __syncwarp();
                if (lane == 0) {
                    s_rec.level = plan.level;
                    s_rec.tree = t;
                    s_rec.node = node - plan.node_base[t];
                    s_rec.split = live ? b.split : -1;
                    s_rec.n_ties = b.ties;
                    s_rec.min_av = b.split >= 0 ? b.av : nan;
                }
                if (lane < kMaxSeg) {
                    s_rec.child[lane] = pop ? rank + first : -1;
                    s_rec.child_pop[lane] = (long long)pop;
                    s_rec.child_stat[lane] = seg ? it_stat[i0 + lane] : 0.0;
                }
                __syncwarp();
                const uint32_t* src = reinterpret_cast<const uint32_t*>(&s_rec);
                uint32_t* dst = reinterpret_cast<uint32_t*>(&plan.records[node]);
                for (int i = lane; i < (int)(sizeof(avt_grow_record) / 4); i += 32) dst[i] = src[i];
            }
        }
    }
    if (lookback && plan.lut2 && node < n_nodes) {   // (look-back mode: the child ids in the row are final)
        __syncthreads();   // warp 0 has written the node's split attribute and bin -> child row
        const int at = plan.node_attr[node];
        const uint16_t* lut = plan.child_lut + (long long)node * plan.lut_w;
        for (int v = tid; v < plan.lut2_J; v += blockDim.x)
            plan.lut2[(long long)node * plan.lut2_J + v] = at >= 0 ? lut[(v >> plan.pk_shift[at]) & plan.pk_mask[at]] : (uint16_t)AVT_NODE_NONE;
    }
    if (lookback) {
        // an empty frontier has no block that owns the next level's node_base: block 0 writes it (all zero)
        if (blockIdx.x == 0 && n_nodes == 0)
            for (int k = tid; k <= plan.n_trees; k += blockDim.x) plan.next_node_base[k] = 0;
        return;
    }
    // ---- last block to finish turns the local child ranks into ids of the next frontier
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        const uint32_t done = atomicAdd(plan.ticket, 1u);
        s_is_last = done == gridDim.x - 1;
        if (s_is_last) {
            *plan.ticket = 0;
            __threadfence();
        }
    }
    __syncthreads();
    if (s_is_last) plan_level_fast(plan, ft.nkids, s_warp_tot, &s_running, s_first);
}

}  // namespace avt
