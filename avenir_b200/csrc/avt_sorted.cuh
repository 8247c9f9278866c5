// Kernel (a) for the deep levels of ONE tree over a wide schema (C5: 512 attributes x 64 bins, frontiers of hundreds to
// thousands of nodes): hist_sorted_kernel, on rows brought into node-contiguous order (hard part 3b).
//
// Why: once a frontier's counters no longer fit a block's shared memory, hist_smem_kernel cannot run at all and
// hist_global_kernel sends one L2 atomic per (row, attribute) -- 10^10 per level of C5, 95 ms per launch on average
// (profiles/bench_r02_c5.json).  With the rows sorted by node id (one 16-bit radix sort of (node id, row) pairs per level,
// cub, plumbing) a block that takes a contiguous span of the order sees only a handful of nodes, so their counters DO fit:
// the block walks its span in windows of S consecutive node ids, counts a window in shared memory and flushes it.
// Counters are laid out [node][class][bin][attribute(32)] as in hist_rot_kernel, but here lane = ATTRIBUTE and a warp
// instruction handles one row: the 32 lanes read the 32 bytes of the row's entry in the rotated row-major copy (one sector,
// lane l takes byte (l - row) mod 32 = attribute l) and issue one conflict-free, fully active ATOMS.  Rows of nodes whose
// counts are derived by sibling subtraction, and rows outside the frontier (they sort to the end), are never touched.
// Included at the end of avt_kernels.cuh.
#pragma once

namespace avt {

constexpr int kSortedThreads = 512;

struct SortedArgs {
    const uint16_t* keys;    // [N] node id of every position of the order, ascending (AVT_NODE_NONE last)
    const uint32_t* perm;    // [N] row at that position
    const uint8_t* cls_s;    // [N] class code at that position
    const uint8_t* w_s;      // [N] multiplicity at that position (HAS_W)
    const uint8_t* rowrot;   // rotated row-major codes, [chunks][rowrot_rows][32]
    long long rowrot_rows;
    const int32_t* bin_off;  // [A+1]
    const uint8_t* skip;     // [n_nodes] 1 = the node's counts are derived, its rows are not counted (nullable)
    uint32_t* H;             // [n_nodes][TB][C]
    long long N, span;       // positions per block
    int A, Bmax, S, TB, C;
};

// order[p] -> class / multiplicity copies in order (two 1-byte gathers per row, once per level instead of once per chunk)
__global__ void sorted_gather_kernel(const uint32_t* perm, const uint8_t* cls, const uint8_t* w, long long N, uint8_t* cls_s, uint8_t* w_s) {
    const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= N) return;
    const uint32_t r = perm[p];
    cls_s[p] = cls[r];
    if (w) w_s[p] = w[r];
}

__global__ void iota_kernel(uint32_t* out, long long N) {
    const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < N) out[p] = (uint32_t)p;
}

template <bool HAS_W>
__global__ void __launch_bounds__(kSortedThreads) hist_sorted_kernel(const SortedArgs a) {
    extern __shared__ __align__(16) uint32_t ssh[];
    constexpr int kWarps = kSortedThreads / 32;
    __shared__ __align__(16) uint8_t s_tile[kWarps * 1024];   // per warp: the 32-byte code entries of the 32 rows it is working on
    __shared__ long long s_wend;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int chunk = blockIdx.y;
    const long long p0 = (long long)blockIdx.x * a.span, p1 = min(a.N, p0 + a.span);
    const uint32_t slot_words = (uint32_t)a.C * a.Bmax * 32u;
    const uint32_t slot_bytes = slot_words * 4u, class_bytes = (uint32_t)a.Bmax * 128u;
    const uint32_t cnt_base = (uint32_t)__cvta_generic_to_shared(ssh);
    const uint32_t tile_base = (uint32_t)__cvta_generic_to_shared(s_tile) + (uint32_t)warp * 1024u;
    const uint32_t lane_base = cnt_base + (uint32_t)lane * 4u;
    const uint8_t* rr = a.rowrot + (long long)chunk * a.rowrot_rows * 32;
    const int at = chunk * 32 + lane;
    const int nbins = at < a.A ? a.bin_off[at + 1] - a.bin_off[at] : 0;
    const int b0 = at < a.A ? a.bin_off[at] : 0;
    for (uint32_t i = threadIdx.x; i < (uint32_t)a.S * slot_words; i += blockDim.x) ssh[i] = 0;   // flushed windows leave zeros behind
    long long cur = p0;
    while (cur < p1) {
        const uint32_t nb = a.keys[cur];              // first node of the window (the same value for every thread)
        if (nb == AVT_NODE_NONE) break;               // rows outside the frontier sort last
        if (threadIdx.x == 0) {
            // end of the window: first position in [cur, p1) whose node id is >= nb + S
            long long lo = cur, hi = p1;
            const uint32_t lim = nb + (uint32_t)a.S;
            while (lo < hi) {
                const long long mid = (lo + hi) >> 1;
                if ((uint32_t)a.keys[mid] < lim) lo = mid + 1;
                else hi = mid;
            }
            s_wend = lo;
        }
        __syncthreads();
        const long long wend = s_wend;
        // Per warp 32 positions at a time.  A position's row entry (32 bytes of the rotated copy = one sector) is fetched by the
        // lane that owns the position and parked in shared memory; the positions of the NEXT round are fetched before this
        // round's updates are issued, so the two dependent loads (order -> row -> codes) are off the critical path.
        uint32_t hb_l = 0xFFFFFFFFu, row_l = 0, w_l = 1;
        uint4 c0 = make_uint4(0, 0, 0, 0), c1 = c0;
        auto fetch = [&](long long q, uint32_t& hb, uint32_t& row, uint32_t& w, uint4& x0, uint4& x1) {
            const long long i = q + lane;
            hb = 0xFFFFFFFFu;
            row = 0;
            w = 1;
            if (i < wend) {
                const uint32_t node = a.keys[i];
                row = a.perm[i];
                if (!(a.skip && a.skip[node])) hb = cnt_base + (node - nb) * slot_bytes + (uint32_t)a.cls_s[i] * class_bytes;
                if (HAS_W) {
                    w = a.w_s[i];
                    if (w == 0u) hb = 0xFFFFFFFFu;
                }
                if (hb != 0xFFFFFFFFu) {   // rows that are not counted (derived node, multiplicity 0) cost no code traffic either
                    const uint4* src = reinterpret_cast<const uint4*>(rr + (long long)row * 32);
                    x0 = __ldg(src);
                    x1 = __ldg(src + 1);
                }
            }
        };
        long long q = cur + (long long)warp * 32;
        if (q < wend) fetch(q, hb_l, row_l, w_l, c0, c1);
        for (; q < wend; q += (long long)kWarps * 32) {
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(tile_base + 32u * lane), "r"(c0.x), "r"(c0.y), "r"(c0.z), "r"(c0.w) : "memory");
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(tile_base + 32u * lane + 16u), "r"(c1.x), "r"(c1.y), "r"(c1.z), "r"(c1.w) : "memory");
            // one shuffle per row: counter block in 128-byte units (16 bits) and the row's rotation (5 bits) in one word
            const uint32_t pk_c = hb_l == 0xFFFFFFFFu ? 0xFFFFFFFFu : ((hb_l - cnt_base) >> 7) | ((row_l & 31u) << 16);
            const uint32_t w_c = w_l;
            __syncwarp();
            const long long qn = q + (long long)kWarps * 32;
            if (qn < wend) fetch(qn, hb_l, row_l, w_l, c0, c1);
#pragma unroll 8
            for (int j = 0; j < 32; ++j) {
                const uint32_t pk = __shfl_sync(0xffffffffu, pk_c, j);
                const uint32_t w = HAS_W ? __shfl_sync(0xffffffffu, w_c, j) : 1u;
                if (pk == 0xFFFFFFFFu) continue;   // warp-uniform: a skipped node, multiplicity 0, or past the window
                uint32_t code;
                asm volatile("ld.shared.u8 %0, [%1];" : "=r"(code) : "r"(tile_base + 32u * j + ((lane - (pk >> 16)) & 31u)));
                red_shared_add(lane_base + ((pk & 0xFFFFu) << 7) + code * 128u, w);
            }
            __syncwarp();
        }
        __syncthreads();
        // flush the window and leave zeros: counter (node, class, bin, attribute) -> H[nb + node][bin_off[attr] + bin][class]
        const uint32_t n_present = (uint32_t)min((long long)a.S, (long long)(a.keys[wend - 1] - nb) + 1);
        for (uint32_t i = threadIdx.x; i < n_present * slot_words; i += blockDim.x) {   // i & 31 == lane
            const uint32_t v = ssh[i];
            if (!v) continue;
            ssh[i] = 0;
            const uint32_t rest = i >> 5;
            const int bin = (int)(rest % (uint32_t)a.Bmax);
            const uint32_t sc = rest / (uint32_t)a.Bmax;
            const int c = (int)(sc % (uint32_t)a.C);
            const uint32_t node = sc / (uint32_t)a.C;
            if (bin >= nbins) continue;
            atomicAdd(a.H + ((long long)(nb + node) * a.TB + b0 + bin) * a.C + c, v);
        }
        __syncthreads();
        cur = wend;
    }
}

}  // namespace avt
