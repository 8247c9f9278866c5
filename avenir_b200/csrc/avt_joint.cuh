// Kernel (a) for narrow schemas whose codes fit one byte per row (call_hangup.json: 1 + 2 + 1 + 4 bits): hist_joint_kernel.
//
// hist_smem_kernel unpacks the packed byte into one code per attribute only to issue one counter update per attribute and
// a byte-permute selection for the fused repartition; ncu (profiles/hist_scan_r01_final_ncu.txt) shows that pass bound by
// instruction issue, not by memory.  Here a block keeps ONE counter space per node indexed by the packed byte itself
// (J = largest packed value + 1 joint codes x classes): a row costs one shared-memory update instead of one per attribute,
// its child under the previous level's split is one 16-bit load from a per-(previous node, packed byte) table, and the
// per-attribute marginals the scan wants are taken once per block at flush time (sum over the highest field first, then
// the low fields from that partial sum).
// Included at the end of avt_kernels.cuh.
#pragma once

namespace avt {

constexpr int kJointThreads = 512;

struct JointArgs {
    const int32_t* bin_off;   // [A+1]
    int A;
    int J;             // joint codes: largest packed value + 1
    int n_prev_max;    // fused repartition: bound on a tree's nodes in the previous frontier (sizes the table)
    int zero_words;    // counter words the block may clear before it waits for the previous launch
    uint32_t shift[4], mask[4];   // field of attribute q in the packed byte
    const uint16_t* lut2_g;       // fused repartition: packed byte -> child rows of the previous frontier, written by the scan kernel
                                  // ([NF_prev][J]); nullptr = derive them here from node_attr / child_lut
};

// Dynamic shared memory: counters [n_nodes + 1][J][C] uint32 (node n_nodes swallows the rows that are not counted), then --
// with the fused repartition -- child table [n_prev + 1][J] uint16 (entry n_prev: every byte -> none).
template <bool HAS_W, bool REPART>
__global__ void __launch_bounds__(kJointThreads) hist_joint_kernel(const HistArgs a, const JointArgs ja) {
    extern __shared__ uint32_t jsh[];
    pdl_launch_dependents();
    for (int i = threadIdx.x; i < ja.zero_words; i += blockDim.x) jsh[i] = 0;
    pdl_wait();   // node_base, the repartition tables and the node ids come from the previous launches
    __syncthreads();
    const int t = blockIdx.x;
    const int nb = a.node_base[t], n_nodes = a.node_base[t + 1] - nb;
    const int J = ja.J, C = a.C;
    const int words = (n_nodes + 1) * J * C;
    for (int i = ja.zero_words + threadIdx.x; i < words; i += blockDim.x) jsh[i] = 0;
    uint16_t* lut2 = reinterpret_cast<uint16_t*>(jsh + words);
    int n_prev = 0;
    if (REPART) {
        const int pnb = a.prev_node_base[t];
        n_prev = min(a.prev_node_base[t + 1] - pnb, ja.n_prev_max);
        if (ja.lut2_g) {
            const uint16_t* src = ja.lut2_g + (long long)pnb * J;
            for (int i = threadIdx.x; i < (n_prev + 1) * J; i += blockDim.x) lut2[i] = i < n_prev * J ? src[i] : (uint16_t)AVT_NODE_NONE;
        } else {
            for (int i = threadIdx.x; i < (n_prev + 1) * J; i += blockDim.x) {
                const int p = i / J, v = i - p * J;
                uint16_t child = (uint16_t)AVT_NODE_NONE;
                if (p < n_prev) {
                    const int at = a.node_attr[pnb + p];
                    if (at >= 0) child = a.child_lut[(long long)(pnb + p) * a.lut_w + ((v >> ja.shift[at]) & ja.mask[at])];
                }
                lut2[i] = child;
            }
        }
    }
    __syncthreads();
    const uint32_t sh_base = (uint32_t)__cvta_generic_to_shared(jsh);
    const uint32_t lut_base = (uint32_t)__cvta_generic_to_shared(lut2);
    const uint16_t* nid = a.nid + (long long)t * a.Npad;
    uint16_t* nid_out = REPART ? a.nid_out + (long long)t * a.Npad : nullptr;
    const bool tree_has_w = HAS_W && a.has_w[t];
    const uint8_t* wt = HAS_W ? a.wts + (long long)t * a.Npad : nullptr;
    const uint32_t JC4 = (uint32_t)(J * C) * 4u, C4 = (uint32_t)C * 4u, J2 = (uint32_t)J * 2u;
    const long long stride = (long long)gridDim.y * blockDim.x * kRowsPerThread;
    for (long long base = ((long long)blockIdx.y * blockDim.x + threadIdx.x) * kRowsPerThread; base < a.N; base += stride) {
        uint4 vn0 = ld_stream(nid + base), vn1 = ld_stream(nid + base + 8);
        const uint4 vp = ld_stream(a.packed + base), vc = ld_stream(a.cls + base);
        uint4 vw = make_uint4(0x01010101u, 0x01010101u, 0x01010101u, 0x01010101u);
        if (tree_has_w) vw = ld_stream(wt + base);
        uint32_t nw[8];
#pragma unroll
        for (int j = 0; j < kRowsPerThread; ++j) {
            const uint32_t p = u4_byte(vp, j);
            uint32_t node = j < 8 ? u4_half(vn0, j) : u4_half(vn1, j - 8);
            if (REPART) {
                asm volatile("ld.shared.u16 %0, [%1];" : "=r"(node) : "r"(lut_base + min(node, (uint32_t)n_prev) * J2 + p * 2u));   // 0xFFFF -> the all-none row
                if (j & 1) nw[j >> 1] |= node << 16;
                else nw[j >> 1] = node;
            }
            const uint32_t w = u4_byte(vw, j);
            const bool dead = node == AVT_NODE_NONE || (HAS_W && w == 0u) || base + j >= a.N;
            const uint32_t slot = dead ? (uint32_t)n_nodes : node;
            red_shared_add(sh_base + slot * JC4 + p * C4 + u4_byte(vc, j) * 4u, HAS_W ? w : 1u);
        }
        if (REPART) {
            *reinterpret_cast<uint4*>(nid_out + base) = make_uint4(nw[0], nw[1], nw[2], nw[3]);
            *reinterpret_cast<uint4*>(nid_out + base + 8) = make_uint4(nw[4], nw[5], nw[6], nw[7]);
        }
    }
    __syncthreads();
    // ---- flush: marginals of every attribute from the joint counters, all threads over (node, output) items.  The highest
    // field (attribute A-1) owns the top bits.  (1) its own marginal: sum over the low part; (2) R[low][class] = sum over the
    // highest field, written over that node's h = 0 counters; (3) every lower attribute's marginal is a sum over R.
    const int top = ja.A - 1;
    const int low_n = 1 << ja.shift[top];                 // values of the low part
    const int high_n = (J + low_n - 1) / low_n;           // values of the highest field that occur
    const int JC = J * C;
    for (int it = threadIdx.x; it < n_nodes * high_n * C; it += blockDim.x) {
        const int node = it / (high_n * C), o = it - node * (high_n * C);
        const int h = o / C, c = o - h * C;
        const uint32_t* cn = jsh + (size_t)node * JC + (size_t)h * low_n * C + c;
        const int lmax = min(low_n, J - h * low_n);
        uint32_t sum = 0;
        for (int l = 0; l < lmax; ++l) sum += cn[l * C];
        if (sum) atomicAdd(a.H + ((long long)(nb + node) * a.TB + ja.bin_off[top] + h) * C + c, sum);
    }
    __syncthreads();
    for (int it = threadIdx.x; it < n_nodes * low_n * C; it += blockDim.x) {
        const int node = it / (low_n * C), o = it - node * (low_n * C);   // o = l * C + c
        uint32_t* cn = jsh + (size_t)node * JC + o;
        uint32_t sum = 0;
        for (int h = 0; h * low_n * C + o < JC; ++h) sum += cn[h * low_n * C];
        cn[0] = sum;   // R lives in the h = 0 counters
    }
    __syncthreads();
    int low_bins = ja.bin_off[top] - ja.bin_off[0];       // bins of the attributes below the highest
    for (int it = threadIdx.x; it < n_nodes * low_bins * C; it += blockDim.x) {
        const int node = it / (low_bins * C), o = it - node * (low_bins * C);
        const int gb = o / C, c = o - gb * C;             // bin index among the low attributes
        int q = 0;
        while (q + 1 < top && gb >= ja.bin_off[q + 1] - ja.bin_off[0]) ++q;
        const int b = gb - (ja.bin_off[q] - ja.bin_off[0]);
        const uint32_t* R = jsh + (size_t)node * JC + c;
        uint32_t sum = 0;
        for (int l = 0; l < low_n; ++l)
            if (((l >> ja.shift[q]) & (int)ja.mask[q]) == b && l * C + c < JC) sum += R[l * C];
        if (sum) atomicAdd(a.H + ((long long)(nb + node) * a.TB + ja.bin_off[q] + b) * C + c, sum);
    }
}

}  // namespace avt
