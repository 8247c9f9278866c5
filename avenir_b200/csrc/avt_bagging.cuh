// Bootstrap multiplicities on the device: the reference's BaggingSampler (explore/BaggingSampler.java:101-122;
// the intended `withReplace` of tree/DecisionTreeBuilder.java:214-229) and the `withoutReplace`
// Bernoulli sub-sampling (tree/DecisionTreeBuilder.java:230-235), driven by a java.util.Random-compatible stream
// (SURVEY App. B-5 / E).  The stream is a 48-bit LCG, so the state before any draw is reachable in O(log k):
// every thread jumps to its own draws and no draw is generated twice.
#pragma once
#include "avt_device.cuh"

namespace avt {

constexpr unsigned long long kLcgMult = 0x5DEECE66DULL;
constexpr unsigned long long kLcgAdd = 0xBULL;
constexpr unsigned long long kLcgMask = (1ULL << 48) - 1;

// k steps of seed' = seed * mult + add (mod 2^48) as one affine map (A, B): seed_k = A * seed + B.
__host__ __device__ inline void lcg_jump(unsigned long long k, unsigned long long& A, unsigned long long& B) {
    unsigned long long a = kLcgMult, c = kLcgAdd;
    A = 1;
    B = 0;
    while (k) {
        if (k & 1) {
            A = A * a;
            B = B * a + c;
        }
        c = (a + 1) * c;
        a = a * a;
        k >>= 1;
    }
    A &= kLcgMask;
    B &= kLcgMask;
}

// java.util.Random.nextDouble() from state s (advanced by two steps)
__device__ __forceinline__ double java_next_double(unsigned long long& s) {
    s = (s * kLcgMult + kLcgAdd) & kLcgMask;
    const unsigned long long hi = s >> (48 - 26);
    s = (s * kLcgMult + kLcgAdd) & kLcgMask;
    const unsigned long long lo = s >> (48 - 27);
    return __dmul_rn((double)((hi << 27) + lo), 1.0 / 9007199254740992.0);
}

struct BagArgs {
    uint8_t* w;               // [T][Npad] multiplicities of the local rows
    long long Npad;
    long long total_rows;     // rows of the whole file (the batches are cut on the file, not on the shard)
    long long row0, n_local;  // the local rows are [row0, row0 + n_local) of the file
    long long seed;           // tree t draws from java.util.Random(seed + t)
    int batch;                // bas.batch.size
    long long first_batch;    // first batch that overlaps the local rows
    int32_t* err;             // set to 1 when a multiplicity exceeds 255
};

// One block per (batch, tree): the batch's m draws are counted in shared memory (m words), then the multiplicities of
// the batch's local rows are written out.  grid = (batches overlapping the local rows, T); dynamic smem = batch * 4.
__global__ void __launch_bounds__(512) bagging_kernel(const BagArgs a) {
    extern __shared__ uint32_t cnt[];
    const long long gb = a.first_batch + blockIdx.x;
    const long long b0 = gb * a.batch;
    const int m = (int)min((long long)a.batch, a.total_rows - b0);
    const int t = blockIdx.y;
    for (int i = threadIdx.x; i < m; i += blockDim.x) cnt[i] = 0;
    __syncthreads();
    // thread i takes the contiguous draws [i * q, (i + 1) * q) of the batch; draw j of the file starts 2 j steps into the stream
    const int q = (m + blockDim.x - 1) / blockDim.x;
    const int d0 = threadIdx.x * q, d1 = min(m, d0 + q);
    if (d0 < d1) {
        unsigned long long A, B;
        lcg_jump(2ULL * (unsigned long long)(b0 + d0), A, B);
        unsigned long long s = ((unsigned long long)(a.seed + t) ^ kLcgMult) & kLcgMask;
        s = (A * s + B) & kLcgMask;
        const double dm = (double)m;
        for (int d = d0; d < d1; ++d) {
            int sel = (int)__dmul_rn(java_next_double(s), dm);
            sel = min(sel, m - 1);
            atomicAdd(&cnt[sel], 1u);
        }
    }
    __syncthreads();
    uint8_t* w = a.w + (long long)t * a.Npad;
    bool over = false;
    for (int i = threadIdx.x; i < m; i += blockDim.x) {
        const long long r = b0 + i - a.row0;
        if (r < 0 || r >= a.n_local) continue;
        const uint32_t c = cnt[i];
        over = over || c > 255u;
        w[r] = (uint8_t)min(c, 255u);
    }
    if (over) *a.err = 1;
}

// Batches too large for shared memory: every draw goes straight to the (zeroed) multiplicity bytes as an atomic add on
// the containing word.  grid = (batches, T, slices of a batch).
__global__ void __launch_bounds__(256) bagging_global_kernel(const BagArgs a) {
    const long long gb = a.first_batch + blockIdx.x;
    const long long b0 = gb * a.batch;
    const long long m = min((long long)a.batch, a.total_rows - b0);
    const int t = blockIdx.y;
    const long long threads = (long long)gridDim.z * blockDim.x, tid = (long long)blockIdx.z * blockDim.x + threadIdx.x;
    const long long q = (m + threads - 1) / threads;
    const long long d0 = tid * q, d1 = min(m, d0 + q);
    if (d0 >= d1) return;
    unsigned long long A, B;
    lcg_jump(2ULL * (unsigned long long)(b0 + d0), A, B);
    unsigned long long s = ((unsigned long long)(a.seed + t) ^ kLcgMult) & kLcgMask;
    s = (A * s + B) & kLcgMask;
    uint32_t* w32 = reinterpret_cast<uint32_t*>(a.w + (long long)t * a.Npad);
    const double dm = (double)m;
    for (long long d = d0; d < d1; ++d) {
        long long sel = (long long)(int)__dmul_rn(java_next_double(s), dm);
        sel = min(sel, m - 1);
        const long long r = b0 + sel - a.row0;
        if (r < 0 || r >= a.n_local) continue;
        const uint32_t sh = 8u * (uint32_t)(r & 3);
        const uint32_t old = atomicAdd(&w32[r >> 2], 1u << sh);
        if (((old >> sh) & 0xFFu) == 0xFFu) *a.err = 1;
    }
}

// `withoutReplace`: row r of the file is kept when (int)(nextDouble() * 100) < rate; one draw per row, in file order.
__global__ void __launch_bounds__(256) bernoulli_kernel(const BagArgs a, int rate, int rows_per_thread) {
    const int t = blockIdx.y;
    const long long l0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * rows_per_thread;
    if (l0 >= a.n_local) return;
    const long long l1 = min(a.n_local, l0 + rows_per_thread);
    unsigned long long A, B;
    lcg_jump(2ULL * (unsigned long long)(a.row0 + l0), A, B);
    unsigned long long s = ((unsigned long long)(a.seed + t) ^ kLcgMult) & kLcgMask;
    s = (A * s + B) & kLcgMask;
    uint8_t* w = a.w + (long long)t * a.Npad;
    for (long long l = l0; l < l1; ++l) w[l] = (int)__dmul_rn(java_next_double(s), 100.0) < rate ? 1 : 0;
}

}  // namespace avt
