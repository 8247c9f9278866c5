// Device-side building blocks shared by the kernels of libavenir_tree.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/avenir_tree.h"
#include "fdlibm_log.cuh"

namespace avt {

constexpr int kMaxC = AVT_MAX_CLASSES;
constexpr int kMaxSeg = AVT_MAX_SEG;

// 128-bit streaming load that does not allocate in L1 (each column byte is used once per pass).
__device__ __forceinline__ uint4 ld_stream(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// shared-memory reduction on a 32-bit shared-window address (no generic-address conversion per update)
__device__ __forceinline__ void red_shared_add(uint32_t addr, uint32_t v) {
    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// programmatic dependent launch: a kernel launched with the programmatic-serialization attribute may start
// while its predecessor in the stream is still running; everything that depends on the predecessor's
// results must come after pdl_wait().  Both are no-ops for ordinary launches.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }

// system-scope accesses for memory shared with other GPUs over NVLink
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// Waits until the flag word another GPU stores into this GPU's memory has reached `epoch` (wrap-safe), but never for
// longer than kPeerWaitNs: a rank that died, returned early or skipped a call must not hang this GPU inside a kernel.
// Returns false on timeout; the caller reports it (error word 3 -> AVT_E_COMM) and carries on with whatever is there.
constexpr unsigned long long kPeerWaitNs = 20ull * 1000 * 1000 * 1000;
__device__ __forceinline__ bool wait_flag_bounded(const uint32_t* flag, uint32_t epoch) {
    if ((int32_t)(ld_acquire_sys(flag) - epoch) >= 0) return true;
    const unsigned long long t0 = global_timer_ns();
    for (;;) {
        for (int i = 0; i < 256; ++i)
            if ((int32_t)(ld_acquire_sys(flag) - epoch) >= 0) return true;
        if (global_timer_ns() - t0 > kPeerWaitNs) return false;
    }
}

__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ uint32_t u4_word(const uint4& v, int i) {
    return i == 0 ? v.x : i == 1 ? v.y : i == 2 ? v.z : v.w;
}
// j-th byte (0..15) of a 16-byte vector
__device__ __forceinline__ uint32_t u4_byte(const uint4& v, int j) {
    return __byte_perm(u4_word(v, j >> 2), 0u, 0x4440u | (uint32_t)(j & 3));   // one PRMT (j is a compile-time constant at every use)
}
// j-th halfword (0..7)
__device__ __forceinline__ uint32_t u4_half(const uint4& v, int j) {
    return __byte_perm(u4_word(v, j >> 1), 0u, (j & 1) ? 0x4432u : 0x4410u);
}

// Impurity of one class-count vector, evaluated exactly as InfoContentStat.processStat
// (util/InfoContentStat.java:71-101): classes with zero count do not exist,
// terms are accumulated in `order`, every operation is a single IEEE rounding (no FMA).
__device__ __forceinline__ double impurity(const uint32_t* cnt, int C, const uint8_t* order,
                                           int criterion, uint32_t total) {
    if (total == 0) return 0.0;
    const double tot = (double)total;
    if (criterion == AVT_CRIT_ENTROPY) {
        const double log2v = 0.6931471805599453;  // StrictMath.log(2)
        double stat = 0.0;
        for (int i = 0; i < C; ++i) {
            uint32_t c = cnt[order[i]];
            if (c == 0) continue;
            double pr = __ddiv_rn((double)c, tot);
            double t = __ddiv_rn(__dmul_rn(pr, strict_log(pr)), log2v);
            stat = __dsub_rn(stat, t);
        }
        return stat;
    }
    double prsq = 0.0;
    for (int i = 0; i < C; ++i) {
        uint32_t c = cnt[order[i]];
        if (c == 0) continue;
        double pr = __ddiv_rn((double)c, tot);
        prsq = __dadd_rn(prsq, __dmul_rn(pr, pr));
    }
    return __dsub_rn(1.0, prsq);
}

}  // namespace avt
