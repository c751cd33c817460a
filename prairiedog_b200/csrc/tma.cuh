// tma.cuh -- bulk asynchronous copies global -> shared memory (TMA, `cp.async.bulk`) completing on an mbarrier.
//
// Used by the partition passes of msd3.cu to fetch the NEXT tile of records while the current one is ranked,
// scattered and written: ncu (profiles/r02_ncu_full_k11_summary.txt) shows passes 2-3 stalled on `long_scoreboard`
// -- a CTA loaded a tile, worked on it, and only then asked for the next one, so few of the resident CTAs had loads
// in flight at any time (DRAM at 46-50 %).  One elected thread issues the copy (no registers, no issue slots of the
// other 255 threads); SASS: UBLKCP + SYNCS.
// PTX forms as in cute/arch/copy_sm90_tma.hpp (SM90_BULK_COPY_G2S) and cutlass/arch/barrier.h (ClusterBarrier).
#pragma once
#include <stdint.h>

namespace pdg {
#ifdef __CUDACC__

__device__ __forceinline__ uint32_t tma_smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// one thread, once; a __syncthreads() must follow before any other thread touches the barrier
__device__ __forceinline__ void tma_bar_init(unsigned long long *bar, uint32_t arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%1], %0;" ::"r"(arrivals), "r"(tma_smem_addr(bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");     // visible to the async proxy
}

// One thread: arms the barrier's current phase with `bytes` and starts the copy.  dst and src 16-byte aligned,
// bytes a multiple of 16 and > 0.  The data is visible to every thread whose tma_bar_wait on that phase returned.
// Every thread must be done READING dst (a barrier since their last read) before this is called.
__device__ __forceinline__ void tma_load(void *dst, const void *src, uint32_t bytes, unsigned long long *bar) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");           // generic-proxy reads of dst before the async write
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%1], %0;" ::"r"(bytes), "r"(tma_smem_addr(bar)) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(tma_smem_addr(dst)), "l"(src), "r"(bytes), "r"(tma_smem_addr(bar))
                 : "memory");
}

// all threads: returns when the phase of the given parity (0, 1, 0, ...) has completed
__device__ __forceinline__ void tma_bar_wait(unsigned long long *bar, uint32_t parity) {
    asm volatile("{\n\t"
                 ".reg .pred P1;\n\t"
                 "LAB_WAIT:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n\t"
                 "@P1 bra DONE;\n\t"
                 "bra LAB_WAIT;\n\t"
                 "DONE:\n\t"
                 "}" ::"r"(tma_smem_addr(bar)), "r"(parity), "r"(0x989680u)
                 : "memory");
}

// The piece of an array of 8-byte records that a bulk copy has to fetch for records [p, p + n): the start moves back
// to a 16-byte boundary (`skip` = 0 or 1 records in front) and the size up to a multiple of 16 (at most one record
// behind) -- the arrays carry that much slack at both ends (msd3_reserve).
struct TmaSpan { const void *src; uint32_t bytes, skip; };
__device__ __forceinline__ TmaSpan tma_span8(const uint2 *p, uint32_t n) {
    const uint32_t skip = (uint32_t)((reinterpret_cast<uintptr_t>(p) >> 3) & 1u);
    return TmaSpan{p - skip, ((n + skip) * 8u + 15u) & ~15u, skip};
}

#endif
}  // namespace pdg
