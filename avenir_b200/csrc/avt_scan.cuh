// Kernel (b): fused impurity scan + per-node arg-min, with the children of the best split and -- in the
// device-driven level loop -- the planning of the next level folded in (last block to finish).
// Included by avt_kernels.cuh after SplitTable / split_av / warp_prefix / ScanBest.
#pragma once

namespace avt {

// Level planning state (device-driven loop): from the scan results derive, per tree, the numbering of the
// children, the bin -> child table and split attribute of every frontier node (inputs of the next fused
// repartition + histogram pass), the next frontier's node_base, and one record per node.
struct PlanArgs {
    const avt_node_best* best;
    const avt_split_eval* best_eval;
    const int32_t* node_base;  // [T+1] current frontier
    int32_t* node_attr;        // [NF]
    uint16_t* child_lut;       // [NF][lut_w]
    int32_t* next_count;       // [T]
    int32_t* prefix;           // [NF + 1] scratch of plan_level: exclusive prefix sum of the nodes' child counts
    avt_grow_record* records;  // [NF] (nullable)
    int32_t* err;              // set to 1 when a tree would exceed the node-id range
    int32_t* next_node_base;   // [T+1] (nullable)
    uint32_t* ticket;          // zero-initialised counter for the last-block election (nullptr: no planning)
    uint32_t* zero_ptr;        // words to clear for the next level's histogram (nullable)
    long long zero_words;
    int lut_w, level, n_trees;
    // peer exchange (rows sharded over several GPUs): the scan block itself sums the node's partial
    // histograms out of every rank's memory over NVLink -- the all-reduce is fused into kernel (b)
    const uint32_t* const* peer_H;  // [n_ranks] this level's partial histogram on every rank (self included); nullptr = single GPU
    const uint32_t* peer_flags;     // [n_ranks] local words; rank r stores its epoch here when its partial histogram is complete
    uint32_t* h_sum;                // local buffer that receives the summed histogram (== the kernel's H argument)
    uint32_t peer_epoch;
    int n_ranks;
    // push form of the peer exchange (small levels, item-parallel scan): each node's block stores its partial counts as
    // (tag << 32 | count) words into every other rank's receive slots and polls its own slots -- local memory only
    unsigned long long* const* push_dst;    // [n_ranks] this rank's slot array inside rank d's memory (nullptr: pull form or single GPU)
    const unsigned long long* push_recv;    // local: slots of source rank s at [s * push_cap, (s + 1) * push_cap)
    const uint32_t* H_local;                // this rank's partial counts of the level
    int push_cap, my_rank;
    uint32_t* const* signal_flags;  // [n_ranks] non-null: THIS kernel tells the ranks that this rank's histogram is complete (its first
                                    // block, right after the wait for the histogram pass) instead of the pass's last block
    // sibling subtraction: the most populous child of every split is NOT counted by the next histogram pass; its
    // counts are derived as parent - (other children), exact in integers, by that child's scan block
    // joint-code histogram kernel (packed narrow schemas): the node's block also expands its bin -> child row into a
    // packed-byte -> child row, so the next histogram pass copies its repartition table instead of deriving it
    uint16_t* lut2;             // [NF][lut2_J] (nullptr: not wanted)
    int lut2_J;
    uint32_t pk_shift[4], pk_mask[4];
    uint8_t* next_skip;         // [NF_next] 1 = derived (nullptr: subtraction off)
    int32_t* next_parent;       // [NF_next] global index (this frontier) of a derived node's parent, else -1
    int32_t* kids;              // [NF][kMaxSeg] global index (next frontier) of each segment's child, -1 = none
    const int32_t* drv_parent;  // [NF] the same table for THIS frontier, written by the previous level's planning (nullptr at level 0)
    const int32_t* prev_kids;   // [NF_prev][kMaxSeg] written by the previous level's planning
    const uint32_t* H_prev;     // complete counts of the previous level
    uint32_t* H_mut;            // this level's counts (the kernel's H argument), completed in place for derived nodes
};

// Tells every rank (itself included) that this rank's partial histogram of epoch `epoch` is complete.
// Launched on the stream right after the histogram pass; remote_flags[r] points INTO rank r's memory.
__global__ void peer_signal_kernel(uint32_t* const* remote_flags, int n_ranks, uint32_t epoch) {
    if ((int)threadIdx.x < n_ranks) {
        __threadfence_system();
        st_release_sys(remote_flags[threadIdx.x], epoch);
    }
}

// Waits until every rank has signalled `epoch` (used once at the end of a level loop, so that no rank
// reuses its exchange buffers while another is still reading them).
__global__ void peer_wait_kernel(const uint32_t* flags, int n_ranks, uint32_t epoch, int32_t* err) {
    if ((int)threadIdx.x < n_ranks)
        if (!wait_flag_bounded(flags + threadIdx.x, epoch)) *err = 3;
}

// The same with a bound on the wait (engine teardown: a rank that is gone must not hang the others).
__global__ void peer_wait_bounded_kernel(const uint32_t* flags, int n_ranks, uint32_t epoch, int max_spins) {
    if ((int)threadIdx.x < n_ranks)
        for (int i = 0; i < max_spins && (int32_t)(ld_acquire_sys(flags + threadIdx.x) - epoch) < 0; ++i) {}
}

// loads that bypass L1: the planning block reads what OTHER blocks of the same launch wrote
__device__ __forceinline__ int32_t ldcg_i32(const int32_t* p) { return __ldcg(p); }
__device__ __forceinline__ uint32_t ldcg_u32(const uint32_t* p) { return __ldcg(p); }
__device__ __forceinline__ long long ldcg_i64(const int64_t* p) { return __ldcg((const long long*)p); }
__device__ __forceinline__ double ldcg_f64(const double* p) { return __ldcg(p); }

// Candidate attributes of the next frontier (getSplitAttributes, tree/DecisionTreeBuilder.java:365-381 with
// tree/SplitManager.java:149-166 `all` and :195-232 `notUsedYet`): a child inherits its parent's candidate set, minus --
// with remove_used -- the attribute its parent was split on (the attributes "on my path" are exactly the ones removed so
// far).  One thread per node of the current frontier; runs behind the scan kernel that planned the level.
__global__ void inherit_masks_kernel(const int32_t* nb_cur, const int32_t* nb_next, int n_trees, const int32_t* node_attr,
                                     const uint16_t* child_lut, int lut_w, const uint32_t* cand_cur, uint32_t* cand_next, int mw,
                                     int remove_used) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= nb_cur[n_trees]) return;
    const int at = node_attr[g];
    if (at < 0) return;
    int lo = 0, hi = n_trees;   // last tree whose first node is <= g (empty trees share their successor's base and are skipped)
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (nb_cur[mid] <= g) lo = mid;
        else hi = mid;
    }
    const int base = nb_next[lo];
    uint32_t last = AVT_NODE_NONE;
    for (int b = 0; b < lut_w; ++b) {
        const uint32_t c = child_lut[(long long)g * lut_w + b];
        if (c == AVT_NODE_NONE || c == last) continue;
        last = c;
        for (int w = 0; w < mw; ++w) {
            uint32_t m = cand_cur[(long long)g * mw + w];
            if (remove_used && w == (at >> 5)) m &= ~(1u << (at & 31));
            cand_next[(long long)(base + (int)c) * mw + w] = m;
        }
    }
}

// Whole-level planning by one block (blockDim.x = kScanThreads), flat over the nodes of ALL trees: a forest level of 100
// trees x a few nodes used to be planned tree by tree, five block barriers each -- 0.6 ms of fixed cost per level of C3, the
// part of the step that does not shrink with more GPUs.  Children are numbered per tree, so with E = exclusive prefix sum of
// the child counts over all nodes (tree-major order), a node's first child is E[g] - E[first node of its tree], the next
// frontier's base of tree t is E[first node of t], and its node count the difference of two such values.
//   pass 1: child counts -> E (block scan, 256 nodes per round, carried in shared memory), kept in p.prefix[NF + 1]
//   pass 2: one thread per node writes its tables, record and skip marks
//   then next_count / next_node_base from E, then (optionally) the zeroing of the next level's histogram region.
__device__ void plan_level(const PlanArgs& p, const SplitTable& st, int* warp_tot, int* running) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int NF = p.node_base[p.n_trees];
    if (threadIdx.x == 0) *running = 0;
    __syncthreads();
    for (int g0 = 0; g0 < NF; g0 += blockDim.x) {
        const int g = g0 + threadIdx.x;
        int nkids = 0;
        if (g < NF) {
            const int best_split = ldcg_i32(&p.best[g].best_split);
            if (ldcg_i64(&p.best[g].population) > 0 && best_split >= 0) {
                const avt_split sp = st.splits[best_split];
                for (int sg = 0; sg < sp.n_seg; ++sg) {
                    uint32_t pop = 0;
                    for (int c = 0; c < st.C; ++c) pop += ldcg_u32(&p.best_eval[g].seg_count[sg][c]);
                    nkids += pop != 0;
                }
            }
        }
        int v = nkids;  // block inclusive scan of nkids
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
        if (g < NF) p.prefix[g] = *running + woff + v - nkids;
        __syncthreads();
        if (threadIdx.x == 0) *running += tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) p.prefix[NF] = *running;
    __syncthreads();   // (one block: its global writes are visible to its own threads behind the barrier)
    for (int g = threadIdx.x; g < NF; g += blockDim.x) {
        int lo = 0, hi = p.n_trees;   // last tree whose first node is <= g (empty trees share their successor's base and are skipped)
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (p.node_base[mid] <= g) lo = mid;
            else hi = mid;
        }
        const int t = lo, nb = p.node_base[t], i = g - nb;
        const int next_base = p.prefix[nb], first = p.prefix[g] - next_base;
        const int best_split = ldcg_i32(&p.best[g].best_split);
        const bool live = ldcg_i64(&p.best[g].population) > 0 && best_split >= 0;
        avt_split sp{};
        if (live) sp = st.splits[best_split];
        p.node_attr[g] = live ? sp.attr : -1;
        uint16_t* lut = p.child_lut + (long long)g * p.lut_w;
        for (int b = 0; b < p.lut_w; ++b) lut[b] = (uint16_t)AVT_NODE_NONE;
        avt_grow_record rec;
        rec.level = p.level; rec.tree = t; rec.node = i;
        rec.split = live ? best_split : -1;
        rec.n_ties = ldcg_i32(&p.best[g].n_ties);
        rec.min_av = ldcg_f64(&p.best[g].min_av);
        for (int sg = 0; sg < kMaxSeg; ++sg) { rec.child[sg] = -1; rec.child_pop[sg] = 0; rec.child_stat[sg] = 0.0; }
        if (live) {
            int k = 0;
            for (int sg = 0; sg < sp.n_seg; ++sg) {
                long long pop = 0;
                for (int c = 0; c < st.C; ++c) pop += ldcg_u32(&p.best_eval[g].seg_count[sg][c]);
                rec.child_pop[sg] = pop;
                rec.child_stat[sg] = ldcg_f64(&p.best_eval[g].seg_stat[sg]);
                if (pop) rec.child[sg] = first + k++;
            }
            if (first + k > (int)AVT_NODE_NONE - 1) *p.err = 1;
            const int B = st.bin_off[sp.attr + 1] - st.bin_off[sp.attr];
            for (int b = 0; b < B; ++b) {
                int sg;
                if (sp.kind == AVT_SPLIT_NUMERIC) {
                    sg = 0;
                    while (sg < sp.n_seg - 1 && b > st.bounds[sp.off + sg]) ++sg;
                } else {
                    sg = st.seg_lut[sp.off + b];
                    if (sg >= sp.n_seg) continue;
                }
                if (rec.child[sg] >= 0) lut[b] = (uint16_t)rec.child[sg];
            }
        }
        if (p.kids) {
            int32_t* kd = p.kids + (long long)g * kMaxSeg;
            int derived = -1;
            if (live && p.next_skip) {   // the most populous child (lowest segment on ties) is derived, not counted
                long long mx = -1;
                for (int sg = 0; sg < sp.n_seg; ++sg)
                    if (rec.child[sg] >= 0 && rec.child_pop[sg] > mx) { mx = rec.child_pop[sg]; derived = sg; }
            }
            for (int sg = 0; sg < kMaxSeg; ++sg) {
                const int gi = rec.child[sg] >= 0 ? next_base + rec.child[sg] : -1;
                kd[sg] = gi;
                if (gi >= 0 && p.next_skip) {
                    p.next_skip[gi] = sg == derived;
                    p.next_parent[gi] = sg == derived ? g : -1;
                }
            }
        }
        if (p.records) p.records[g] = rec;
    }
    for (int t = threadIdx.x; t <= p.n_trees; t += blockDim.x) {
        const int e0 = p.prefix[p.node_base[t]];
        if (t < p.n_trees) p.next_count[t] = p.prefix[p.node_base[t + 1]] - e0;
        if (p.next_node_base) p.next_node_base[t] = e0;
    }
    if (p.zero_ptr)
        for (long long i = threadIdx.x; i < p.zero_words; i += blockDim.x) p.zero_ptr[i] = 0;
}

// One block per frontier node (blocks beyond the true frontier only take part in the last-block
// election).  Warps take the node's candidate attributes round-robin; a warp stages the attribute's
// prefix counts in shared memory, its lanes evaluate the attribute's splits and write av[node][split];
// the block reduces to (min av, lowest split index, number of ties).  Meanwhile the last warp computes
// the node's own class counts and impurity; afterwards warp 0 computes the children of the best split,
// one lane per segment.
// The scan of ONE node by one block of kScanThreads threads (see scan_kernel).
__device__ __forceinline__ void scan_node(const uint32_t* H, const SplitTable& st, const uint32_t* cand_mask, const uint8_t* order_g,
                                          int criterion, double* av_buf, avt_node_best* out, avt_split_eval* best_eval, int node,
                                          const PlanArgs& plan) {
    __shared__ uint32_t Pbuf[kScanWarps][kScanWordsPerWarp];
    __shared__ ScanBest wbest[kScanWarps];
    __shared__ uint8_t order[kMaxC];
    __shared__ uint32_t s_cnt[kMaxC];
    __shared__ double s_node_stat;
    __shared__ uint32_t s_pop;
    __shared__ avt_split_eval s_ev;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int C = st.C;
    {
        if (threadIdx.x < C) order[threadIdx.x] = order_g[threadIdx.x];
        __syncthreads();
        if (plan.peer_H) {
            // every rank's partial counts for this node, pulled over NVLink and summed; the ranks' frontiers are
            // identical, so node indices agree.  Wait first until all of them have finished their histogram pass.
            if ((int)threadIdx.x < plan.n_ranks)
                if (!wait_flag_bounded(plan.peer_flags + threadIdx.x, plan.peer_epoch)) *plan.err = 3;
            __syncthreads();
            const long long words = (long long)st.TB * C, off = (long long)node * words;
            const int dp = plan.drv_parent ? plan.drv_parent[node] : -1;
            auto pull = [&](long long o, long long i) {
                uint32_t sum = 0;
                for (int r0 = 0; r0 < plan.n_ranks; r0 += 8) {   // 8 NVLink loads in flight per thread
                    uint32_t v[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) v[k] = r0 + k < plan.n_ranks ? ld_relaxed_sys(plan.peer_H[r0 + k] + o + i) : 0u;
#pragma unroll
                    for (int k = 0; k < 8; ++k) sum += v[k];
                }
                return sum;
            };
            for (long long i = threadIdx.x; i < words; i += blockDim.x) {
                uint32_t sum;
                if (dp < 0) {
                    sum = pull(off, i);
                } else {   // derived node: parent's complete counts minus the counted siblings' (summed over the ranks)
                    sum = plan.H_prev[(long long)dp * words + i];
                    for (int k = 0; k < kMaxSeg; ++k) {
                        const int kid = plan.prev_kids[(long long)dp * kMaxSeg + k];
                        if (kid >= 0 && kid != node) sum -= pull((long long)kid * words, i);
                    }
                }
                plan.h_sum[off + i] = sum;
            }
            __syncthreads();
        } else if (plan.drv_parent) {
            const int dp = plan.drv_parent[node];
            if (dp >= 0) {   // derived node: complete its (uncounted) slots in place
                const long long words = (long long)st.TB * C;
                for (long long i = threadIdx.x; i < words; i += blockDim.x) {
                    uint32_t v = plan.H_prev[(long long)dp * words + i];
                    for (int k = 0; k < kMaxSeg; ++k) {
                        const int kid = plan.prev_kids[(long long)dp * kMaxSeg + k];
                        if (kid >= 0 && kid != node) v -= plan.H_mut[(long long)kid * words + i];
                    }
                    plan.H_mut[(long long)node * words + i] = v;
                }
            }
            __syncthreads();
        }
        const uint32_t* Hnode = H + (long long)node * st.TB * C;
        const int mask_words = (st.A + 31) >> 5;
        const double nan = __longlong_as_double(0x7ff8000000000000LL);
        if (warp == kScanWarps - 1) {
            // the node's own class counts (bins of attribute 0 partition its rows) and impurity
            if (lane < C) {
                const int B0 = st.bin_off[1] - st.bin_off[0];
                uint32_t s = 0;
                for (int bb = 0; bb < B0; ++bb) s += Hnode[bb * C + lane];
                s_cnt[lane] = s;
            }
            __syncwarp();
            if (lane == 0) {
                uint32_t n = 0;
                for (int c = 0; c < C; ++c) n += s_cnt[c];
                s_pop = n;
                s_node_stat = impurity(s_cnt, C, order, criterion, n);
            }
        }
        ScanBest best{0.0, -1, 0};
        for (int at = warp; at < st.A; at += kScanWarps) {
            const int s0 = st.attr_first[at], s1 = st.attr_first[at + 1];
            if (s0 == s1) continue;
            const bool cand = cand_mask == nullptr || ((cand_mask[(long long)node * mask_words + (at >> 5)] >> (at & 31)) & 1u);
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
        if (warp == 0) {
            ScanBest b = wbest[0];
            for (int w = 1; w < kScanWarps; ++w) best_merge(b, wbest[w]);
            const uint32_t n = s_pop;
            if (best_eval) {
                // children of the (lowest-index) best split: one lane per segment, then lane 0 accumulates
                // the weighted average in segment order exactly as split_av does
                uint32_t* evw = reinterpret_cast<uint32_t*>(&s_ev);
                for (int i = lane; i < (int)(sizeof(avt_split_eval) / 4); i += 32) evw[i] = 0;
                __syncwarp();
                const bool have = b.split >= 0 && n > 0;
                avt_split sp{};
                if (have) sp = st.splits[b.split];
                if (have && lane < sp.n_seg) {
                    const int Bs = st.bin_off[sp.attr + 1] - st.bin_off[sp.attr];
                    uint32_t cnt[kMaxC];
                    segment_counts(st, sp, lane, Bs, nullptr, Hnode + (long long)st.bin_off[sp.attr] * C, cnt);
                    uint32_t m = 0;
                    for (int c = 0; c < C; ++c) {
                        m += cnt[c];
                        s_ev.seg_count[lane][c] = cnt[c];
                    }
                    s_ev.seg_stat[lane] = m ? impurity(cnt, C, order, criterion, m) : 0.0;
                }
                __syncwarp();
                if (lane == 0) {
                    double av = nan;
                    if (have) {
                        double weighted = 0.0;
                        uint32_t total = 0;
                        for (int j = 0; j < sp.n_seg; ++j) {
                            uint32_t m = 0;
                            for (int c = 0; c < C; ++c) m += s_ev.seg_count[j][c];
                            if (m) {
                                weighted = __dadd_rn(weighted, __dmul_rn(s_ev.seg_stat[j], (double)m));
                                total += m;
                            }
                        }
                        av = __ddiv_rn(weighted, (double)total);
                    }
                    s_ev.av = av;
                }
                __syncwarp();
                uint32_t* dst = reinterpret_cast<uint32_t*>(&best_eval[node]);
                for (int i = lane; i < (int)(sizeof(avt_split_eval) / 4); i += 32) dst[i] = evw[i];
            }
            if (lane == 0) {
                avt_node_best r;
                r.min_av = b.split >= 0 ? b.av : nan;
                r.node_stat = s_node_stat;
                r.population = n;
                r.best_split = b.split;
                r.n_ties = b.ties;
                out[node] = r;
            }
        }
    }
}

__global__ void __launch_bounds__(kScanThreads) scan_kernel(const uint32_t* H, SplitTable st, const uint32_t* cand_mask,
                                                            const uint8_t* order_g, int criterion, double* av_buf,
                                                            avt_node_best* out, avt_split_eval* best_eval,
                                                            const int32_t* n_nodes_dev, PlanArgs plan) {
    __shared__ int s_warp_tot[kScanWarps];
    __shared__ int s_running;
    __shared__ int s_is_last;
    pdl_launch_dependents();
    pdl_wait();   // the histogram (and the frontier size) come from the previous launch
    if (plan.signal_flags && blockIdx.x == 0 && (int)threadIdx.x < plan.n_ranks)
        st_release_sys(plan.signal_flags[threadIdx.x], plan.peer_epoch);   // the pass is complete: its counts are in this GPU's L2
    const int node = blockIdx.x;
    if (!(n_nodes_dev && node >= *n_nodes_dev)) scan_node(H, st, cand_mask, order_g, criterion, av_buf, out, best_eval, node, plan);
    if (plan.ticket == nullptr) return;
    // ---- last block to finish plans the next level
    __threadfence();  // every thread publishes its results before the block takes its ticket
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const uint32_t done = atomicAdd(plan.ticket, 1u);
        s_is_last = done == gridDim.x - 1;
        if (s_is_last) {
            *plan.ticket = 0;
            __threadfence();
        }
    }
    __syncthreads();
    if (s_is_last) plan_level(plan, st, s_warp_tot, &s_running);
}

// Stand-alone planning launch (one block), used by the read-back form of the device-driven loop.
__global__ void __launch_bounds__(kScanThreads) plan_kernel(PlanArgs p, SplitTable st) {
    __shared__ int s_warp_tot[kScanWarps];
    __shared__ int s_running;
    plan_level(p, st, s_warp_tot, &s_running);
}

}  // namespace avt
