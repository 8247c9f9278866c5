// Kernel (a) for bagged forests and wide schemas: hist_forest_kernel.
//
// Why a second histogram kernel: on forest shapes (C3: 100 trees x 32 attributes) hist_smem_kernel is bound by the
// shared-memory pipe, not by HBM -- ncu shows l1tex at 99 % with 3.5-3.9 wavefronts per ATOMS (32 lanes = 32 rows hitting
// random bins collide on banks), every tree re-reads all bin codes, and rows with multiplicity 0 (37 % of a bootstrap
// sample) or in nodes derived by sibling subtraction still issue their updates (a partially active ATOMS runs at a
// fraction of the full-warp rate on sm_100a, so predicating them away is slower still; tools/microbench/atoms_bench.cu).
//
// Here one warp instruction updates ONE row: lane = attribute, and the block's private counters are laid out
// [tree][slot][class][bin][attribute(32)], so the bank of an update is its lane -- every ATOMS is a single conflict-free
// wavefront -- and rows that are not counted cost no update at all (no partial warps: a row is either skipped by the
// whole warp or updated by the whole warp).  A warp owns 32-row tiles: lane a keeps column a's 32 codes of the tile in
// registers (pre-scaled to counter offsets) and loops over the trees resident in the block, so the codes are read once per
// (tile, tree group) instead of once per tree; per (tile, tree) only 64 B of node ids and 32 B of multiplicities move,
// prefetched through a per-warp cp.async ring.  The previous level's repartition is folded in as in hist_smem_kernel.
#pragma once
// (included at the end of avt_kernels.cuh)

namespace avt {

constexpr int kForestThreads = 512;
constexpr int kForestWarps = kForestThreads / 32;
constexpr int kForestRing = 8;          // (tile, tree) units in flight per warp
constexpr int kForestUnitBytes = 96;    // 32 node ids (uint16) + 32 multiplicities (uint8)

struct ForestArgs {
    const int32_t* bin_off;   // [A+1]
    int A;
    int G;             // trees resident per block (grid.x = ceil(T / G))
    int n_slots;       // counter slots (counted nodes) per tree in shared memory
    int Bmax;          // bins per attribute in the shared layout (max over all attributes)
    int n_nodes_max;   // bound on a tree's frontier nodes (stride of the node -> slot tables)
    long long tiles_per_part;   // 32-row tiles per grid.y slice
    int32_t* err;      // set to 4 when a tree has more counted nodes than n_slots (host bound violated)
};

__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src, bool valid) {
    const int sz = valid ? 4 : 0;   // src-size 0: the four bytes are zero-filled
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
// explicit shared-window accesses for the hot loop (a generic pointer into dynamic shared memory compiles to LD.E/ST.E)
__device__ __forceinline__ uint32_t lds_u16(uint32_t addr) {
    uint16_t v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts_u32(uint32_t addr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Shared memory of a block (dynamic):
//   counters   [G][n_slots][C][Bmax][32] uint32
//   slot_of    [G][n_nodes_max + 1] uint16      node -> slot, 0xFFFF = not counted here
//   node_of    [G][n_slots] uint16              slot -> node (flush)
//   ring       [warps][kForestRing][96 B]       prefetched node ids + multiplicities
//   bcast      [warps][32] uint32               per-row (counter base | multiplicity << 20), read back as broadcasts
//   has_w      [G] uint8
template <bool HAS_W, bool REPART>
__global__ void __launch_bounds__(kForestThreads, 1) hist_forest_kernel(const HistArgs a, const ForestArgs f) {
    extern __shared__ __align__(16) uint32_t fsh[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int t0 = blockIdx.x * f.G;
    const int Gb = min(f.G, a.T - t0);
    const int chunk = blockIdx.z;                 // attributes [32 * chunk, 32 * chunk + 32)
    const int attr = chunk * 32 + lane;
    const bool attr_ok = attr < f.A;
    const uint32_t slot_words = (uint32_t)a.C * f.Bmax * 32u;       // words of one (tree, slot)
    const uint32_t cnt_words = (uint32_t)f.G * f.n_slots * slot_words;
    uint32_t* cnt = fsh;
    uint16_t* slot_of = reinterpret_cast<uint16_t*>(cnt + cnt_words);
    const int so_stride = f.n_nodes_max + 1;
    uint16_t* node_of = slot_of + (size_t)f.G * so_stride;
    uint8_t* ring_all = reinterpret_cast<uint8_t*>(node_of + (size_t)f.G * f.n_slots);
    ring_all = reinterpret_cast<uint8_t*>(((uintptr_t)ring_all + 15) & ~(uintptr_t)15);
    uint8_t* ring = ring_all + (size_t)warp * kForestRing * kForestUnitBytes;
    uint32_t* bcast = reinterpret_cast<uint32_t*>(ring_all + (size_t)kForestWarps * kForestRing * kForestUnitBytes) + warp * 32;
    uint8_t* s_has_w = reinterpret_cast<uint8_t*>(bcast - warp * 32 + kForestWarps * 32);   // [G]

    pdl_launch_dependents();
    for (uint32_t i = threadIdx.x; i < cnt_words; i += blockDim.x) cnt[i] = 0;
    pdl_wait();   // node_base, skip marks, repartition tables and node ids come from the previous launches
    // node -> slot tables: the counted (not derived) nodes of each resident tree, numbered in node order
    for (int g = warp; g < Gb; g += kForestWarps) {
        const int nb = a.node_base[t0 + g], n = a.node_base[t0 + g + 1] - nb;
        for (int i = lane; i < so_stride; i += 32) slot_of[g * so_stride + i] = 0xFFFF;
        if (lane == 0) s_has_w[g] = HAS_W && a.has_w[t0 + g] ? 1 : 0;
        __syncwarp();
        int run = 0;
        for (int i0 = 0; i0 < n; i0 += 32) {
            const int i = i0 + lane;
            const bool counted = i < n && !(a.skip && a.skip[nb + i]);
            const uint32_t m = __ballot_sync(0xffffffffu, counted);
            const int s = run + __popc(m & ((1u << lane) - 1u));
            if (i < n && i < so_stride) slot_of[g * so_stride + i] = (counted && s < f.n_slots) ? (uint16_t)s : (uint16_t)0xFFFF;
            if (counted && s < f.n_slots) node_of[g * f.n_slots + s] = (uint16_t)i;
            run += __popc(m);
        }
        if (lane == 0 && (run > f.n_slots || n > f.n_nodes_max)) *f.err = 4;
        for (int s = run + lane; s < f.n_slots; s += 32) node_of[g * f.n_slots + s] = 0xFFFF;
    }
    __syncthreads();

    const uint32_t cnt_base = (uint32_t)__cvta_generic_to_shared(cnt);
    const uint32_t ring_base = (uint32_t)__cvta_generic_to_shared(ring);
    const uint32_t bcast_base = (uint32_t)__cvta_generic_to_shared(bcast);
    const uint32_t slot_of_base = (uint32_t)__cvta_generic_to_shared(slot_of);
    const uint32_t has_w_base = (uint32_t)__cvta_generic_to_shared(s_has_w);
    const uint8_t* col = attr_ok ? (const uint8_t*)a.cols[attr] : nullptr;
    const long long n_tiles = (a.N + 31) >> 5;
    const long long tile_lo = (long long)blockIdx.y * f.tiles_per_part;
    const long long tile_hi = min(n_tiles, tile_lo + f.tiles_per_part);
    // this warp's tiles: tile_lo + warp, + kForestWarps, ...
    const long long my_tiles = tile_hi > tile_lo + warp ? (tile_hi - tile_lo - warp + kForestWarps - 1) / kForestWarps : 0;
    const uint32_t class_bytes = (uint32_t)f.Bmax * 128u;          // one (slot, class): Bmax bins x 32 attributes x 4 B
    const uint32_t slot_bytes = slot_words * 4u;
    const uint32_t tree_bytes = (uint32_t)f.n_slots * slot_bytes;

    static_assert((kForestRing & (kForestRing - 1)) == 0, "ring size must be a power of two");
    auto tile_of = [&](int k) { return tile_lo + warp + (long long)k * kForestWarps; };
    // prefetch of one unit's node ids (lanes 0-15: two rows each) and multiplicities (lanes 16-23: four rows each);
    // units are (tile k, tree g) in the order the loop below consumes them
    int ik = 0, ig = 0;
    uint32_t iu = 0;
    auto issue_next = [&]() {
        if (ik < (int)my_tiles) {
            const long long row0 = tile_of(ik) << 5;
            const uint32_t dst = ring_base + (iu & (kForestRing - 1)) * kForestUnitBytes;
            if (lane < 16) {
                const long long r = row0 + 2 * lane;
                cp_async4(dst + 4u * lane, a.nid + (long long)(t0 + ig) * a.Npad + r, r < a.Npad);
            } else if (HAS_W && lane < 24) {
                const long long r = row0 + 4 * (lane - 16);
                const bool hw = lds_u8(has_w_base + ig) != 0;
                cp_async4(dst + 64u + 4u * (lane - 16), a.wts + (long long)(t0 + ig) * a.Npad + r, hw && r < a.Npad);
            }
        }
        cp_async_commit();
        ++iu;
        if (++ig == Gb) { ig = 0; ++ik; }
    };
    auto load_codes = [&](long long tile, uint4& c0, uint4& c1, uint32_t& cl) {
        const long long row0 = tile << 5;
        c0 = make_uint4(0, 0, 0, 0);
        c1 = c0;
        if (attr_ok) {
            c0 = ld_stream(col + row0);
            if (row0 + 16 < a.Npad) c1 = ld_stream(col + row0 + 16);
        }
        cl = row0 + lane < a.N ? a.cls[row0 + lane] : 0u;
    };

    for (int d = 0; d < kForestRing - 1; ++d) issue_next();
    uint4 nc0 = make_uint4(0, 0, 0, 0), nc1 = nc0;
    uint32_t ncl = 0;
    if (my_tiles) load_codes(tile_of(0), nc0, nc1, ncl);
    uint32_t cu = 0;
    for (int k = 0; k < (int)my_tiles; ++k) {
        const long long row = (tile_of(k) << 5) + lane;
        // new tile: codes -> counter offsets (row j of the tile: code * 128 + lane * 4, bin-major / attribute-minor, packed in
        // pairs), then start loading the next tile's codes
        uint32_t offs[16];
        {
            const uint32_t l4 = 4u * lane;
            const uint32_t w8[8] = {nc0.x, nc0.y, nc0.z, nc0.w, nc1.x, nc1.y, nc1.z, nc1.w};
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const uint32_t b0 = w8[q] & 0xFFu, b1 = (w8[q] >> 8) & 0xFFu, b2 = (w8[q] >> 16) & 0xFFu, b3 = w8[q] >> 24;
                offs[2 * q] = (b0 * 128u + l4) | ((b1 * 128u + l4) << 16);
                offs[2 * q + 1] = (b2 * 128u + l4) | ((b3 * 128u + l4) << 16);
            }
        }
        const uint32_t cl = ncl;
        if (k + 1 < (int)my_tiles) load_codes(tile_of(k + 1), nc0, nc1, ncl);
        for (int g = 0; g < Gb; ++g, ++cu) {
            issue_next();
            cp_async_wait<kForestRing - 1>();   // this unit has landed (a thread's groups complete in order)
            __syncwarp();
            const uint32_t slot = ring_base + (cu & (kForestRing - 1)) * kForestUnitBytes;
            uint32_t node = lds_u16(slot + 2u * lane);
            uint32_t w = 1u;
            if (HAS_W && lds_u8(has_w_base + g)) w = lds_u8(slot + 64u + lane);
            if (row >= a.N) node = AVT_NODE_NONE;
            const int t = t0 + g;
            if (REPART) {
                uint32_t nn = AVT_NODE_NONE;
                if (node != AVT_NODE_NONE) {
                    const int gi = a.prev_node_base[t] + (int)node;
                    const int at = __ldg(a.node_attr + gi);
                    if (at >= 0) {
                        const uint32_t code = __ldg((const uint8_t*)a.cols[at] + row);
                        nn = __ldg(a.child_lut + (long long)gi * a.lut_w + code);
                    }
                }
                node = nn;
                if (chunk == 0 && row < a.Npad) a.nid_out[(long long)t * a.Npad + row] = (uint16_t)node;
            }
            uint32_t s = 0xFFFFu;
            if (node != AVT_NODE_NONE && w != 0u) s = lds_u16(slot_of_base + 2u * (g * so_stride + min(node, (uint32_t)f.n_nodes_max)));
            const bool live = s != 0xFFFFu;
            const uint32_t m = __ballot_sync(0xffffffffu, live);
            if (m) {
                sts_u32(bcast_base + 4u * lane, (cnt_base + (uint32_t)g * tree_bytes + s * slot_bytes + cl * class_bytes) | (w << 20));
                __syncwarp();
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    if ((m >> (4 * q)) & 0xFu) {
                        const uint4 v = lds_v4(bcast_base + 16u * q);   // same address in every lane: one broadcast
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            const int j = 4 * q + r;
                            if ((m >> j) & 1u) {
                                const uint32_t x = u4_word(v, r);
                                const uint32_t off = (j & 1) ? (offs[j >> 1] >> 16) : (offs[j >> 1] & 0xFFFFu);
                                red_shared_add((x & 0xFFFFFu) + off, HAS_W ? (x >> 20) : 1u);
                            }
                        }
                    }
                }
            }
            __syncwarp();   // bcast and the ring slot are rewritten by the next iterations
        }
    }
    cp_async_wait<0>();
    __syncthreads();
    // flush: counters [g][slot][class][bin][lane] -> H[node][bin_off[attr] + bin][class]
    for (int g = 0; g < Gb; ++g) {
        const int nb = a.node_base[t0 + g];
        const uint32_t per_tree = (uint32_t)f.n_slots * slot_words;
        for (uint32_t i = threadIdx.x; i < per_tree; i += blockDim.x) {
            const uint32_t v = cnt[(uint32_t)g * per_tree + i];
            if (!v) continue;
            const int la = (int)(i & 31u);
            const uint32_t rest = i >> 5;
            const int bin = (int)(rest % (uint32_t)f.Bmax);
            const uint32_t sc = rest / (uint32_t)f.Bmax;
            const int c = (int)(sc % (uint32_t)a.C);
            const int s = (int)(sc / (uint32_t)a.C);
            const int at = chunk * 32 + la;
            if (at >= f.A) continue;   // lanes without an attribute counted into their own (unused) column
            const int node = node_of[g * f.n_slots + s];
            const int b0 = f.bin_off[at];
            if (node == 0xFFFF || bin >= f.bin_off[at + 1] - b0) continue;
            atomicAdd(a.H + ((long long)(nb + node) * a.TB + b0 + bin) * a.C + c, v);
        }
    }
}

}  // namespace avt
