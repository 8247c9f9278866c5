// Developer microbenchmark: throughput of red.shared.add on sm_100a under the access patterns the forest histogram
// pass can choose between.  Prints updates per clock per SM.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void red32(uint32_t addr, uint32_t v) { asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ void red64(uint32_t addr, unsigned long long v) { asm volatile("red.shared.add.u64 [%0], %1;" ::"r"(addr), "l"(v) : "memory"); }

// mode 0: random words in `words`; 1: conflict-free (lane-distinct banks); 2: random, odd lanes predicated off;
// 3: random, upper half-warps off (compacted would be: half the warps); 4: u64 random; 5: LDS+ADD+STS (non-atomic, racy) random
// 6: random within 16-word groups chosen by lane (attribute-per-lane layout)
template <int MODE>
__global__ void __launch_bounds__(512) k(uint32_t* out, int iters, uint32_t words, long long* cyc) {
    extern __shared__ uint32_t sh[];
    for (uint32_t i = threadIdx.x; i < words; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(sh);
    uint32_t a[16];
    uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    const uint32_t lane = threadIdx.x & 31;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        s = s * 1664525u + 1013904223u;
        uint32_t w = (s >> 8) % words;
        if (MODE == 1) w = (w & ~31u) | lane;
        if (MODE == 4) w &= ~1u;
        if (MODE == 6) w = ((w & ~511u) | (lane << 4) | (w & 15u)) % words;
        a[j] = w * 4u;
    }
    const uint32_t step = (MODE == 1) ? 128u * 37u : (MODE == 4 ? 8u * 1031u : (MODE == 6 ? 2048u * 3u : 4u * 1031u));
    const uint32_t bytes = words * 4u;
    long long t0 = clock64();
    const bool on = MODE == 2 ? !(lane & 1) : (MODE == 3 ? lane < 16 : true);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            if (on) {
                if (MODE == 4) red64(base + a[j], 1ull);
                else if (MODE == 5) { uint32_t* p = (uint32_t*)((char*)sh + a[j]); *(volatile uint32_t*)p = *(volatile uint32_t*)p + 1; }
                else red32(base + a[j], 1u);
            }
            a[j] += step;
            if (a[j] >= bytes) a[j] -= bytes;
        }
    }
    __syncthreads();
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    uint32_t acc = 0;
    for (uint32_t i = threadIdx.x; i < words; i += blockDim.x) acc += sh[i];
    if (acc == 0xdeadbeef) out[0] = acc;
}

template <int MODE>
void run(const char* name, int threads, int blocks_per_sm, uint32_t words, int sms) {
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 4); cudaMalloc(&cyc, 8 * sms * blocks_per_sm);
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    const int iters = 2000;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<sms * blocks_per_sm, threads, words * 4>>>(out, 10, words, cyc);
    cudaEventRecord(e0);
    k<MODE><<<sms * blocks_per_sm, threads, words * 4>>>(out, iters, words, cyc);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h[2048]; cudaMemcpy(h, cyc, 8 * sms * blocks_per_sm, cudaMemcpyDeviceToHost);
    double mean = 0; for (int i = 0; i < sms * blocks_per_sm; ++i) mean += h[i]; mean /= sms * blocks_per_sm;
    double frac = MODE == 2 || MODE == 3 ? 0.5 : 1.0;
    double upd = (double)iters * 16 * threads * blocks_per_sm * frac;
    printf("%-34s thr=%4d blk/SM=%d words=%6u : %.2f updates/clk/SM (in-kernel clocks), %.3f ms, %.2f T updates/s %s\n", name, threads, blocks_per_sm, words,
           upd / mean, ms, upd * sms / (ms * 1e-3) / 1e12, err == cudaSuccess ? "" : cudaGetErrorString(err));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    printf("%s, %d SMs, %d MHz\n", p.name, sms, p.clockRate / 1000);
    for (int bps = 1; bps <= 2; ++bps) {
        run<0>("red.u32 random", 512, bps, 16384, sms);
        run<0>("red.u32 random", 512, bps, 1664, sms);
        run<0>("red.u32 random (big)", 512, bps, 26000, sms);
        run<1>("red.u32 conflict-free", 512, bps, 16384, sms);
        run<2>("red.u32 random, odd lanes off", 512, bps, 16384, sms);
        run<3>("red.u32 random, half-warp off", 512, bps, 16384, sms);
        run<4>("red.u64 random", 512, bps, 16384, sms);
        run<5>("lds+add+sts random (racy)", 512, bps, 16384, sms);
        run<6>("red.u32 lane=attr, 16-bin groups", 512, bps, 16384, sms);
    }
    run<0>("red.u32 random", 256, 1, 16384, sms);
    run<0>("red.u32 random", 128, 1, 16384, sms);
    run<0>("red.u32 random", 1024, 1, 16384, sms);
    run<0>("red.u32 random 64 words", 512, 2, 64, sms);
    return 0;
}
