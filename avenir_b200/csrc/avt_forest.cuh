// Kernel (a) for bagged forests and wide schemas: hist_rot_kernel (+ rowrot_kernel, which builds its input layout).
//
// Why a second histogram kernel: on forest shapes (C3: 100 trees x 32 attributes) hist_smem_kernel is bound by the
// shared-memory pipe, not by HBM -- ncu (profiles/hist_c3_r02_baseline_ncu.txt) shows l1tex at 99 % with 3.5-3.9 wavefronts
// per ATOMS (32 lanes = 32 rows hitting random bins collide on banks) -- and every tree re-reads and re-unpacks all bin codes.
// A first redesign with lane = attribute (one row per warp instruction, rows that are not counted skipped by whole warps,
// live (row, tree) pairs handed from a per-lane "row" phase to a per-lane "attribute" phase through shared-memory lists) was
// conflict-free too but spent 190-290 warp instructions per (32-row tile, tree) on bookkeeping and ran at 1.6 ms (16 trees
// resident) to 8.5 ms (1 tree resident) per 10 M rows x 16 trees, against 2.3-2.6 ms for hist_smem_kernel; measured and
// dropped (profiles/README.md).  hist_rot_kernel below needs 1.4-2.0 ms at every depth.
#pragma once
// (included at the end of avt_kernels.cuh)

namespace avt {

constexpr int kForestThreads = 512;
constexpr int kForestWarps = kForestThreads / 32;

struct ForestArgs {
    const int32_t* bin_off;   // [A+1]
    int A;
    int G;             // trees resident per block (grid.x = ceil(T / G))
    int n_slots;       // counter slots (counted nodes) per tree in shared memory
    int Bmax;          // bins per attribute in the shared layout (max over all attributes)
    int n_nodes_max;   // bound on a tree's frontier nodes (stride of the node -> slot tables)
    int n_prev_max;    // fused repartition: bound on a tree's nodes in the previous frontier
    long long tiles_per_part;   // 32-row tiles per grid.y slice
    int32_t* err;      // set to 4 when a tree has more counted nodes than n_slots (host bound violated)
    const uint8_t* rowrot;      // hist_rot_kernel: rotated row-major copy of the codes, [chunks][rowrot_rows][32]
    long long rowrot_rows;      // rows per chunk in `rowrot` (padded row count + one tile of slack)
    int pf;            // node ids / multiplicities are pulled into L2 this many trees ahead of their use (0 = off)
};

// explicit shared-window accesses for the hot loops (a generic pointer into dynamic shared memory compiles to LD.E/ST.E)
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
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void sts_v4(uint32_t addr, const uint4& v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ----------------------------------------------------------------------------------------------------------------------
// hist_rot_kernel: lane = ROW, attributes visited in a lane-rotated order.
//
// The block's private counters are laid out [tree][slot][class][bin][attribute(32)]: the bank of a counter is its attribute.
// Every lane owns one row of the warp's 32-row tile and, at step p of the 32-step update loop, counts attribute
// (p + lane) mod 32 of its row: at every step the 32 lanes touch 32 different attributes, i.e. 32 different banks -- each
// ATOMS is one conflict-free wavefront with all lanes active, three instructions per counter update, nothing shuffled or
// broadcast.  For that a lane must find "attribute (p + lane) mod 32 of my row" at a compile-time byte position p of its
// registers: the data set keeps a second, row-major copy of the codes in which row r's 32 bytes are rotated by r mod 32
// (rowrot_kernel, built once when the data set is installed), so one 32-byte load per row and tile gives exactly that.
// The tile's codes stay in registers (pre-scaled to counter offsets) across the loop over the block's resident trees; per
// (row, tree) only the node id, the multiplicity, the fused repartition (shared-memory tables) and the slot lookup remain.
// Rows that are not counted (outside the frontier, multiplicity 0, node derived by sibling subtraction) update a phantom
// slot: a partially active ATOMS is much slower than a full one on sm_100a (tools/microbench/atoms_bench.cu).
// Shared memory: counters [G * n_slots + 1 phantom][C][Bmax][32], tile copies [warps][32 rows][32 B] (fused repartition),
// then lut32 [G][n_prev_max][lut_w] (bin -> slot << 16 | child of every previous node), slot_of [G][n_nodes_max + 1] (node -> slot,
// 0xFFFF = not counted), node_of [G][n_slots], attr16 [G][n_prev_max] (split attribute, 0xFFFF = not expanded), has_w [G].
// Every global row buffer (columns, classes, node ids, multiplicities, the rotated copy) is allocated with 32 rows of slack, so
// a tile that reaches past the padded row count reads in bounds; such rows are never counted and never stored.
__global__ void rowrot_kernel(const void* const* cols, int A, long long N, long long rows_alloc, uint8_t* out) {
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int chunk = blockIdx.y;
    if (r >= rows_alloc) return;
    uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (r < N) {
        const int rot = (int)(r & 31);
#pragma unroll 4
        for (int p = 0; p < 32; ++p) {
            const int a = chunk * 32 + ((p + rot) & 31);
            const uint32_t code = a < A ? ((const uint8_t*)cols[a])[r] : 0u;
            w[p >> 2] |= code << (8 * (p & 3));
        }
    }
    uint4* dst = reinterpret_cast<uint4*>(out + ((long long)chunk * rows_alloc + r) * 32);
    dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
    dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
}

template <bool HAS_W, bool REPART>
__global__ void __launch_bounds__(kForestThreads, 1) hist_rot_kernel(const HistArgs a, const ForestArgs f) {
    extern __shared__ __align__(16) uint32_t fsh[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int t0 = blockIdx.x * f.G;
    const int Gb = min(f.G, a.T - t0);
    const int chunk = blockIdx.z;
    const uint32_t slot_words = (uint32_t)a.C * f.Bmax * 32u;
    const uint32_t cnt_words = ((uint32_t)f.G * f.n_slots + 1u) * slot_words;    // + the phantom slot
    const int so_stride = f.n_nodes_max + 1;
    const int lut_stride = f.n_prev_max * a.lut_w;
    uint32_t* cnt = fsh;
    uint8_t* codes_s = reinterpret_cast<uint8_t*>(cnt + cnt_words);
    uint32_t* lut32 = reinterpret_cast<uint32_t*>(codes_s + (REPART ? kForestWarps * 1024 : 0));
    uint16_t* slot_of = reinterpret_cast<uint16_t*>(lut32 + (REPART ? (size_t)f.G * lut_stride : 0));
    uint16_t* node_of = slot_of + (size_t)f.G * so_stride;
    uint16_t* attr16 = node_of + (size_t)f.G * f.n_slots;   // split attribute of every previous node, 0xFFFF = not expanded
    uint8_t* s_has_w = reinterpret_cast<uint8_t*>(attr16 + (REPART ? (size_t)f.G * f.n_prev_max : 0));

    pdl_launch_dependents();
    for (uint32_t i = threadIdx.x; i < cnt_words; i += blockDim.x) cnt[i] = 0;
    pdl_wait();
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
            if (i < n && i < so_stride && counted && s < f.n_slots) slot_of[g * so_stride + i] = (uint16_t)s;
            if (counted && s < f.n_slots) node_of[g * f.n_slots + s] = (uint16_t)i;
            run += __popc(m);
        }
        if (lane == 0 && (run > f.n_slots || n > f.n_nodes_max)) *f.err = 4;
        for (int s = run + lane; s < f.n_slots; s += 32) node_of[g * f.n_slots + s] = 0xFFFF;
    }
    __syncthreads();
    if (REPART) {
        for (int g = 0; g < Gb; ++g) {
            const int pnb = a.prev_node_base[t0 + g], n_prev = min(a.prev_node_base[t0 + g + 1] - pnb, f.n_prev_max);
            if (threadIdx.x == 0 && a.prev_node_base[t0 + g + 1] - pnb > f.n_prev_max) *f.err = 4;
            for (int i = threadIdx.x; i < n_prev * a.lut_w; i += blockDim.x) {
                const uint32_t child = a.child_lut[(long long)pnb * a.lut_w + i];
                const uint32_t s = child == AVT_NODE_NONE ? 0xFFFFu : (uint32_t)slot_of[g * so_stride + min(child, (uint32_t)f.n_nodes_max)];
                lut32[g * lut_stride + i] = (s << 16) | child;
            }
            for (int i = threadIdx.x; i < n_prev; i += blockDim.x) {
                const int at = a.node_attr[pnb + i];
                attr16[g * f.n_prev_max + i] = at < 0 ? (uint16_t)0xFFFF : (uint16_t)at;
            }
        }
        __syncthreads();
    }
    uint32_t cnt_base = (uint32_t)__cvta_generic_to_shared(cnt);
    uint32_t codes_base = (uint32_t)__cvta_generic_to_shared(codes_s) + (uint32_t)warp * 1024u + 32u * lane;   // this lane's row
    uint32_t lut_base = (uint32_t)__cvta_generic_to_shared(lut32);
    uint32_t slot_of_base = (uint32_t)__cvta_generic_to_shared(slot_of);
    uint32_t attr16_base = (uint32_t)__cvta_generic_to_shared(attr16);
    uint32_t has_w_base = (uint32_t)__cvta_generic_to_shared(s_has_w);
    asm volatile("" : "+r"(cnt_base), "+r"(codes_base), "+r"(lut_base), "+r"(slot_of_base), "+r"(attr16_base), "+r"(has_w_base));
    const uint32_t class_bytes = (uint32_t)f.Bmax * 128u;
    const uint32_t slot_bytes = slot_words * 4u;
    const uint32_t tree_bytes = (uint32_t)f.n_slots * slot_bytes;
    const uint32_t phantom = cnt_base + (uint32_t)f.G * tree_bytes;   // class 0 of the phantom slot: any bin, own column
    const long long n_tiles = (a.N + 31) >> 5;
    const long long tile_lo = (long long)blockIdx.y * f.tiles_per_part;
    const long long tile_hi = min(n_tiles, tile_lo + f.tiles_per_part);
    const uint8_t* rr = f.rowrot + (long long)chunk * f.rowrot_rows * 32;
    // column (= attribute within the chunk) this lane counts at step p, as a byte offset: 4 * ((p + lane) & 31), two per word
    uint32_t colw[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) colw[q] = (4u * ((2u * q + lane) & 31u)) | ((4u * ((2u * q + 1u + lane) & 31u)) << 16);

    uint4 nc0 = make_uint4(0, 0, 0, 0), nc1 = nc0;
    uint32_t ncl = 0;
    auto load_row = [&](long long tile) {
        const long long row = (tile << 5) + lane;
        const uint4* src = reinterpret_cast<const uint4*>(rr + row * 32);
        nc0 = __ldg(src);
        nc1 = __ldg(src + 1);
        ncl = a.cls[row];
    };
    const uint32_t npad32 = (uint32_t)a.Npad;
    const int pf_tiles = f.pf > 0 ? max(1, (f.pf + Gb - 1) / max(Gb, 1)) : 0;   // at least one tile, at least `pf` trees ahead
    long long tile = tile_lo + warp;
    if (tile < tile_hi) load_row(tile);
    for (; tile < tile_hi; tile += kForestWarps) {
        const long long row = (tile << 5) + lane;
        const bool row_ok = row < a.N;
        uint32_t offs[16];   // step p: code * 128 + column offset, two steps per register
        {
            const uint32_t w8[8] = {nc0.x, nc0.y, nc0.z, nc0.w, nc1.x, nc1.y, nc1.z, nc1.w};
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                offs[2 * q] = __byte_perm(w8[q], 0u, 0x4140u) * 128u + colw[2 * q];
                offs[2 * q + 1] = __byte_perm(w8[q], 0u, 0x4342u) * 128u + colw[2 * q + 1];
            }
            if (REPART) {
                sts_v4(codes_base, nc0);          // own row, own 32 bytes: no other lane reads or writes them
                sts_v4(codes_base + 16u, nc1);
            }
        }
        const uint32_t cl = ncl;
        if (tile + kForestWarps < tile_hi) load_row(tile + kForestWarps);
        const uint16_t* nid_row = a.nid + (long long)t0 * a.Npad + row;
        uint16_t* nid_out_row = REPART ? a.nid_out + (long long)t0 * a.Npad + row : nullptr;
        const uint8_t* w_row = HAS_W ? a.wts + (long long)t0 * a.Npad + row : nullptr;
        // node id and multiplicity of this row in the next two trees are in flight while a tree's updates issue
        auto fetch = [&](int g, uint32_t& n, uint32_t& w) {
            n = AVT_NODE_NONE;
            w = 1u;
            if (g < Gb) {
                n = nid_row[(uint32_t)g * npad32];   // 32-bit element index: the host keeps G * Npad below 2^32 (one IMAD.WIDE per address)
                if (HAS_W && lds_u8(has_w_base + g)) w = w_row[(uint32_t)g * npad32];
            }
        };
        // L2 prefetch of the 64 + 32 bytes a (tile, tree) pair reads, one whole tile ahead: when the warp starts a tile it requests,
        // for every tree of the block, the node ids / multiplicities of the tile it will visit `pf_tiles` rounds from now (the
        // two-tree register prefetch covers one DRAM latency at best; measured on C3 at 20 M rows x 100 trees: 140.7 -> 125.0 ms per
        // step).  Lanes 0 and 16 only: the two 32-byte sectors of a tile's node ids.
        if (f.pf > 0 && (lane & 15) == 0) {
            const long long ft = tile + (long long)pf_tiles * kForestWarps;
            if (ft < tile_hi) {
                const uint16_t* pn = a.nid + (long long)t0 * a.Npad + (ft << 5) + lane;
                const uint8_t* pw = HAS_W ? a.wts + (long long)t0 * a.Npad + (ft << 5) : nullptr;
                for (int g = 0; g < Gb; ++g) {
                    prefetch_l2(pn + (uint32_t)g * npad32);
                    if (HAS_W && lane == 0) prefetch_l2(pw + (uint32_t)g * npad32);
                }
            }
        }
        uint32_t n0, w0, n1, w1;
        fetch(0, n0, w0);
        fetch(1, n1, w1);
        for (int g = 0; g < Gb; ++g) {
            uint32_t node = row_ok ? n0 : AVT_NODE_NONE;
            const uint32_t w = w0;
            n0 = n1;
            w0 = w1;
            fetch(g + 2, n1, w1);
            uint32_t s = 0xFFFFu;
            if (REPART) {
                uint32_t e = 0xFFFFFFFFu;
                if (node != AVT_NODE_NONE) {
                    node = min(node, (uint32_t)f.n_prev_max - 1u);
                    const uint32_t at = lds_u16(attr16_base + 2u * (g * f.n_prev_max + node));
                    if (at != 0xFFFFu) {
                        // byte position of attribute `at` in this row's rotated copy; another chunk's column comes from global memory
                        const uint32_t code = (at >> 5) == (uint32_t)chunk ? lds_u8(codes_base + ((at - (uint32_t)lane) & 31u))
                                                                           : (uint32_t)__ldg((const uint8_t*)a.cols[at] + row);
                        e = lds_u32(lut_base + 4u * (g * lut_stride + node * a.lut_w + code));
                    }
                }
                node = e & 0xFFFFu;
                s = e >> 16;
                if (chunk == 0 && row < a.Npad) nid_out_row[(uint32_t)g * npad32] = (uint16_t)node;
            } else if (node != AVT_NODE_NONE) {
                s = lds_u16(slot_of_base + 2u * (g * so_stride + min(node, (uint32_t)f.n_nodes_max)));
            }
            const bool live = s != 0xFFFFu && w != 0u;
            const uint32_t hb = live ? cnt_base + (uint32_t)g * tree_bytes + s * slot_bytes + cl * class_bytes : phantom;
#pragma unroll
            for (int q = 0; q < 16; ++q) {
                red_shared_add(hb + (offs[q] & 0xFFFFu), HAS_W ? w : 1u);
                red_shared_add(hb + (offs[q] >> 16), HAS_W ? w : 1u);
            }
        }
    }
    __syncthreads();
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
            if (at >= f.A) continue;
            const int node = node_of[g * f.n_slots + s];
            const int b0 = f.bin_off[at];
            if (node == 0xFFFF || bin >= f.bin_off[at + 1] - b0) continue;
            atomicAdd(a.H + ((long long)(nb + node) * a.TB + b0 + bin) * a.C + c, v);
        }
    }
}

// ----------------------------------------------------------------------------------------------------------------------
// repartition_rot_kernel: the stand-alone repartition (kernel (c), last level of avt_grow_levels) for forests over <= 32
// uint8 attributes.  repartition_kernel gathers one code byte per (row, tree) from the column of that tree's split attribute:
// a 32-byte sector per byte, 320 GB of sectors for C3's 10^10 pairs (39 ms per call, 6 % of the step).  Here lane = row, the
// row's 32 codes come once from the rotated row-major copy (one sector per row, parked in shared memory) and the loop runs
// over ALL trees: per (row, tree) only the node id is read and written.  grid.x row parts, 512 threads.
__global__ void __launch_bounds__(kForestThreads) repartition_rot_kernel(uint16_t* nid_all, const uint8_t* rowrot, const int32_t* node_base,
                                                                          const int32_t* node_attr, const uint16_t* child_lut, int lut_w, int T,
                                                                          long long N, long long Npad, long long tiles_per_part) {
    __shared__ __align__(16) uint8_t codes_s[kForestWarps * 1024];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t codes_base = (uint32_t)__cvta_generic_to_shared(codes_s) + (uint32_t)warp * 1024u + 32u * lane;
    const long long n_tiles = (N + 31) >> 5;
    const long long tile_lo = (long long)blockIdx.x * tiles_per_part, tile_hi = min(n_tiles, tile_lo + tiles_per_part);
    for (long long tile = tile_lo + warp; tile < tile_hi; tile += kForestWarps) {
        const long long row = (tile << 5) + lane;
        const uint4* src = reinterpret_cast<const uint4*>(rowrot + row * 32);
        sts_v4(codes_base, __ldg(src));            // own row, own 32 bytes: no other lane touches them
        sts_v4(codes_base + 16u, __ldg(src + 1));
        uint16_t* nid = nid_all + row;
        if ((lane & 15) == 0)
            for (int t = 0; t < min(T, 8); ++t) prefetch_l2(nid + (long long)t * Npad);
        uint32_t next = row < Npad ? nid[0] : AVT_NODE_NONE;
        for (int t = 0; t < T; ++t) {
            const uint32_t old = next;
            if (t + 1 < T && row < Npad) next = nid[(long long)(t + 1) * Npad];
            if ((lane & 15) == 0 && t + 8 < T) prefetch_l2(nid + (long long)(t + 8) * Npad);
            uint32_t nw = AVT_NODE_NONE;
            if (old != AVT_NODE_NONE && row < N) {
                const int g = node_base[t] + (int)old;
                const int at = node_attr[g];
                if (at >= 0) nw = child_lut[(long long)g * lut_w + lds_u8(codes_base + (((uint32_t)at - (uint32_t)lane) & 31u))];
            }
            if (row < Npad) nid[(long long)t * Npad] = (uint16_t)nw;
        }
    }
}

}  // namespace avt
