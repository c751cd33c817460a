// msd3.cuh -- constants and the per-tile partition core of the offsets-first MSD sort (msd3.cu).
#pragma once
#include "traverse.cuh"

namespace pdg {

#ifndef M3_CFG_THREADS
#define M3_CFG_THREADS 256   /* tools/tune_msd3.py: 256 threads x 8 items, 5 CTAs/SM was the fastest of 8 shapes on B200 */
#define M3_CFG_ITEMS 8
#define M3_CFG_MINB 5
#endif
static constexpr int M3_THREADS = M3_CFG_THREADS;
static constexpr int M3_WARPS = M3_THREADS / 32;
static constexpr int M3_ITEMS = M3_CFG_ITEMS;
static constexpr int M3_MINB = M3_CFG_MINB;                       // CTAs per SM the pass kernels are compiled for
static constexpr int M3_TILE = M3_THREADS * M3_ITEMS;   // items per tile
static constexpr int M3_WARP_TILE = 32 * M3_ITEMS;      // consecutive items per warp
static constexpr int M3_MAX_BINS = 256;
static constexpr int M3_BPT = (M3_MAX_BINS + M3_THREADS - 1) / M3_THREADS;   // bins a thread owns in the scan
static constexpr int M3_TILE_WORDS = M3_TILE / 32;      // packed words per position tile
static constexpr int kMsd3MaxBits = 22;                 // k <= 11: digits (<=7, <=7, <=8)
static constexpr uint32_t M3_SIMPLE = 0x80000000u;      // tile_info flag: every window is a k-mer

// dynamic shared memory of one tile partition
struct M3Smem {
    uint2 stage[M3_TILE];                    // the tile in sorted order: {position, code}
    uint32_t wcnt[M3_WARPS][M3_MAX_BINS];    // per (warp, bin): running count, then base inside the tile
    uint32_t delta[M3_MAX_BINS];             // global offset of the bin's run minus its start in stage
    uint32_t cursor[M3_MAX_BINS];            // global offset where this CTA writes the bin next
    uint32_t wsum[M3_WARPS];
};

#ifdef __CUDACC__
// k-mer code of position (32 * word + lane): hi / lo are the packed words `word` and `word + 1`
// (big endian, 32 residues each).  Works on 32-bit halves: one funnel shift per window.
__device__ __forceinline__ uint32_t m3_code_at(uint2 hi, uint2 lo, int lane, int k) {
    const uint32_t x = lane < 16 ? hi.y : hi.x, y = lane < 16 ? hi.x : lo.y;
    return __funnelshift_l(y, x, 2 * (lane & 15)) >> (32 - 2 * k);
}

// Validity bits of the 32 windows starting in packed word `word` (positions 32*word ..): bit l is
// set iff some contig holds the whole window.  ci: any contig index with begin[ci] <= 32*word, or 0.
__device__ __forceinline__ uint32_t m3_valid_word(const ContigTable &ct, uint64_t word, uint64_t ci, int k) {
    const uint64_t p0 = word << 5, p1 = p0 + 32;
    uint32_t m = 0;
    while (ci + 1 < ct.n && ct.begin[ci + 1] <= p0) ++ci;
    for (; ci < ct.n; ++ci) {
        const uint64_t b = ct.begin[ci];
        if (b >= p1) break;
        const uint64_t e = ct.end[ci];
        if (e < p0 + k) continue;
        const uint64_t lo = b > p0 ? b - p0 : 0, hi = e - k - p0;   // valid starts [lo, hi] relative to p0
        if (lo > 31) continue;
        const uint32_t upto = hi >= 31 ? 0xffffffffu : ((2u << (uint32_t)hi) - 1u);
        m |= upto & ~((1u << (uint32_t)lo) - 1u);
    }
    return m;
}

// Stable partition of one tile by a digit of NBITS bits.  Item j of lane l of warp w is tile
// element w*32*M3_ITEMS + j*32 + l; dg[j] in [0, BINS) or BINS (= not a record, dropped; dg is clobbered).
// rec(j) returns the record {position, code} of item j (called once, for kept items).  On entry sm.cursor[b] is the global offset of the first record
// of bin b of this tile; on exit it has advanced by the tile's count (a CTA can walk consecutive
// tiles of a range).  OUT4: write the position to out4, else (code << 32 | position) to out8.
// Ranking: the lanes of a round that share a digit find each other with NBITS ballots (no shared
// memory, no MATCH.ANY -- ~0.5 lanes/clk/SM on sm_100); the per-(warp, bin) running count is one
// 32-bit word read by all and advanced by the lowest lane of the group.  ncu on the earlier
// {mask, count} table ranked by a shared-memory atomicOr: 22 of 37 shared-memory wavefronts per 32
// records went into ranking (ATOMS.OR alone 9); this form needs ~7.
// Needs: wcnt zeroed and a barrier since; leaves wcnt zeroed and ends with a barrier.
// One digit bit of the ranking ballots: lanes whose bit equals mine stay in `same` / never enter `other`.
__device__ __forceinline__ void m3_vote_bit(uint32_t d, uint32_t bitmask, uint32_t &same, uint32_t &other) {
    asm("{\n\t.reg .pred p;\n\t.reg .b32 m, t;\n\t"
        "and.b32 t, %2, %3;\n\t"
        "setp.ne.u32 p, t, 0;\n\t"
        "vote.sync.ballot.b32 m, p, 0xffffffff;\n\t"
        "@p and.b32 %0, %0, m;\n\t"
        "@!p or.b32 %1, %1, m;\n\t}"
        : "+r"(same), "+r"(other) : "r"(d), "r"(bitmask));
}

template <bool OUT4, int NBITS, typename RecFn, typename DigitFn>
__device__ __forceinline__ void m3_partition_tile(M3Smem &sm, RecFn rec, uint32_t (&dg)[M3_ITEMS],
                                                  const int BINS, DigitFn digit, uint64_t *__restrict__ out8,
                                                  uint32_t *__restrict__ out4) {
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t *wc = sm.wcnt[w];
    // ballots and the running count of the (warp, bin) pair, item by item; dg[j] becomes digit | rank << 16
    // (ranking all items first and walking the counts afterwards cost ~10% more: the packed
    // intermediate had to be unpacked again)
#pragma unroll
    for (int j = 0; j < M3_ITEMS; ++j) {
        const uint32_t d = dg[j], dd = d & (M3_MAX_BINS - 1);
        const bool ok = d != (uint32_t)BINS;
        uint32_t same = __ballot_sync(0xffffffffu, ok), other = 0u;
#pragma unroll
        for (int b = 0; b < NBITS; ++b) m3_vote_bit(d, 1u << b, same, other);
        const uint32_t peers = same & ~other, below = peers & lt;
        const uint32_t cnt = wc[dd];
        dg[j] = d | ((cnt + __popc(below)) << 16);
        __syncwarp();
        if (ok && below == 0u) wc[dd] = cnt + __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // bins tid*BPT .. tid*BPT+BPT-1: counts of the warps -> exclusive bases, tile count
    uint32_t c[M3_BPT][M3_WARPS], tot[M3_BPT], sum = 0;
#pragma unroll
    for (int i = 0; i < M3_BPT; ++i) {
        const int b = tid * M3_BPT + i;
        tot[i] = 0;
        if (b < BINS) {
#pragma unroll
            for (int ww = 0; ww < M3_WARPS; ++ww) { c[i][ww] = sm.wcnt[ww][b]; tot[i] += c[i][ww]; }
        }
        sum += tot[i];
    }
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) sm.wsum[w] = inc;
    __syncthreads();
    uint32_t base = 0, all = 0;
#pragma unroll
    for (int ww = 0; ww < M3_WARPS; ++ww) {
        const uint32_t v = sm.wsum[ww];
        if (ww < w) base += v;
        all += v;
    }
    uint32_t run = base + inc - sum;               // start of this thread's first bin inside the tile
#pragma unroll
    for (int i = 0; i < M3_BPT; ++i) {
        const int b = tid * M3_BPT + i;
        if (b < BINS) {
            const uint32_t cur = sm.cursor[b];
            sm.delta[b] = cur - run;
            sm.cursor[b] = cur + tot[i];
#ifdef M3_REREAD_COUNTS
            // variant for tools/tune_msd3.py ("reread"): read the per-warp counts again instead of keeping them
            // in registers across the scan (passes 2-3 spill them: 32 / 40 B of stack at 48 registers).
            // Measured on B200 at 5e8 records, passes 1 / 2 / 3: 2.19 / 2.39 / 2.79 ms against 2.15 / 2.25 / 2.88
            // -- helps pass 3 only, so it stays off until the choice is per pass.
#pragma unroll
            for (int ww = 0; ww < M3_WARPS; ++ww) { const uint32_t cw = sm.wcnt[ww][b]; sm.wcnt[ww][b] = run; run += cw; }
#else
#pragma unroll
            for (int ww = 0; ww < M3_WARPS; ++ww) { sm.wcnt[ww][b] = run; run += c[i][ww]; }
#endif
        }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < M3_ITEMS; ++j) {
        const uint32_t d = dg[j] & 0xffffu;
        if (d != (uint32_t)BINS) sm.stage[wc[d] + (dg[j] >> 16)] = rec(j);
    }
    __syncthreads();
    // re-arm the table for the next tile while the stage drains
    for (int i = tid; i < M3_WARPS * M3_MAX_BINS; i += M3_THREADS) (&sm.wcnt[0][0])[i] = 0u;
#pragma unroll
    for (int j = 0; j < M3_ITEMS; ++j) {
        const uint32_t s = j * M3_THREADS + tid;
        if (s < all) {
            const uint2 x = sm.stage[s];
            const uint32_t at = sm.delta[digit(x.y)] + s;
            if (OUT4) out4[at] = x.x; else out8[at] = ((uint64_t)x.y << 32) | x.x;
        }
    }
    __syncthreads();
}

__device__ __forceinline__ void m3_clear_table(M3Smem &sm) {
    for (int i = threadIdx.x; i < M3_WARPS * M3_MAX_BINS; i += M3_THREADS) (&sm.wcnt[0][0])[i] = 0u;
}
#endif

}  // namespace pdg
