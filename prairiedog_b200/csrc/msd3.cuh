// msd3.cuh -- constants and the per-tile partition core of the offsets-first MSD sort (msd3.cu).
#pragma once
#include "traverse.cuh"
#include "tma.cuh"

namespace pdg {

#ifndef M3_CFG_THREADS
#define M3_CFG_THREADS 256   /* tools/tune_msd3.py: 256 threads x 8 items, 5 CTAs/SM was the fastest of 8 shapes on B200 */
#define M3_CFG_ITEMS 8
#define M3_CFG_MINB 5
#endif
static constexpr int M3_THREADS = M3_CFG_THREADS;
static constexpr int M3_WARPS = M3_THREADS / 32;
static constexpr int M3_ITEMS = M3_CFG_ITEMS;
static constexpr int M3_MINB = M3_CFG_MINB;                       // CTAs per SM passes 1-2 are compiled for
#ifndef M3_CFG_MINB3
#define M3_CFG_MINB3 4   /* pass 3 is the one pass that spills at 48 registers: 4 CTAs/SM (64 registers) 2.69 ms, 5: 2.81 ms */
#endif
static constexpr int M3_MINB3 = M3_CFG_MINB3;
// widths of the low and the middle digit at k = 11 (the high digit takes the rest, <= 7 bits).  8|7: buckets of pass 3
// of ~30 K records at config 3; 7|8: ~15 K records that stay in L2 between pass 3's two sweeps -- measured: pass 3
// 2.64 -> 2.43 ms, pass 2 2.14 -> 2.83 ms (64-byte runs per bin and tile), profiles/r02_tune_split78.txt.  8|7 it is.
#ifndef M3_CFG_B1
#define M3_CFG_B1 8
#define M3_CFG_B2 7
#endif
static constexpr int M3_B1 = M3_CFG_B1, M3_B2 = M3_CFG_B2;
static_assert(M3_B1 <= 8 && M3_B2 <= 8 && M3_B1 + M3_B2 == 15, "digit split");
static constexpr int M3_TILE = M3_THREADS * M3_ITEMS;   // items per tile
static constexpr int M3_WARP_TILE = 32 * M3_ITEMS;      // consecutive items per warp
static constexpr int M3_MAX_BINS = 256;
static constexpr int M3_BPT = (M3_MAX_BINS + M3_THREADS - 1) / M3_THREADS;   // bins a thread owns in the scan
static constexpr int M3_TILE_WORDS = M3_TILE / 32;      // packed words per position tile
static constexpr int kMsd3MaxBits = 22;                 // k <= 11: digits (<=7, <=7, <=8)
static constexpr uint32_t M3_SIMPLE = 0x80000000u;      // tile_info flag: every window is a k-mer

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

// Stable partition of one tile by a digit of NBITS bits (m3_partition_tile2 below).  Item j of lane l of warp w is
// tile element w*32*M3_ITEMS + j*32 + l; dg[j] in [0, BINS) or BINS (= not a record, dropped).  rec(j) returns the
// record {position, code} of item j (called once, for kept items).  On entry sm.cursor[b] is the global offset of
// the first record of bin b of this tile; on exit it has advanced by the tile's count (a CTA can walk consecutive
// tiles of a range).  OUT4: write the position to out4, else (code << 32 | position) to out8.
// Ranking: the lanes of a round that share a digit find each other with NBITS ballots (no shared memory, no
// MATCH.ANY -- ~137 clk per warp instruction on B200, profiles/r02_pipelab.txt); the per-(warp, bin) running count is
// one 32-bit word read by all and advanced by the lowest lane of the group.
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

// ---- v2 of the tile partition (round 2) ---------------------------------------------------------------------
// Same algorithm and the same results as m3_partition_tile; what changed is the instruction count.  Measured on
// B200 (tools/pipelab.cu, profiles/r02_pipelab.txt): one 7-bit ranking round costs ~26 clk per SMSP, i.e. only
// ~0.35 ms of a 2.2-2.8 ms pass -- the passes were bound by the OTHER ~60-100 instructions per record round
// (SASS of round 1: 92 per item in pass 2: ranking 52 of which ~30 needed, bin scan 14, scatter 9, write-out 16).
//   * FULL = every item of the tile is a record: no validity ballot, no guards, no sentinel digit;
//   * the (warp, bin) slot found while ranking is kept (packed with the rank) and reused by the scatter;
//   * the bin scan is done by BINS threads only;
//   * the write-out of a FULL tile is straight-line code (no per-item branch).
// Measured on B200 at config 3 (profiles/r02_tune_core_v2_v3.txt, r02_tune_core_v4.txt): against round 1's core this
// one executes 21 % fewer instructions per item and is NOT faster (pass 1 2.15 -> 2.04, pass 2 2.23 -> 2.32, pass 3
// 2.80 -> 2.81 ms); a variant with the counts as bytes (one word per bank: no bank conflict on any count access) was
// slower still (pass 2 2.53 ms).  Neither instruction issue nor count-table conflicts bound these passes.
// The count table stays row-major, wcnt[warp][bin] (bank = bin mod 32: the lanes of a warp spread over all banks).
// A bin-major table would let the scan move a bin's 8 counts as two uint4, but then the lanes of one warp would
// only ever touch 4 banks ((8 bin + warp) mod 32) in the ranking loop, the hottest one -- rejected on paper.
struct M3Smem2 {
    uint2 stage[M3_TILE];
    uint32_t wcnt[M3_WARPS * M3_MAX_BINS];   // [warp][bin]: running count, then base inside the tile
    uint32_t delta[M3_MAX_BINS];
    uint32_t cursor[M3_MAX_BINS];
    uint32_t wsum[M3_WARPS];
};
__device__ __forceinline__ void m3_clear_table2(M3Smem2 &sm) {
    uint4 *t = reinterpret_cast<uint4 *>(sm.wcnt);
#pragma unroll
    for (int i = threadIdx.x; i < M3_WARPS * M3_MAX_BINS / 4; i += M3_THREADS) t[i] = make_uint4(0u, 0u, 0u, 0u);
}

// after_rank(): called by every thread right after the barrier that ends the ranking -- the first point at which no
// thread needs the tile's INPUT any more (the records are in registers); passes 2-3 start the copy of the next tile there
struct M3NoHook { __device__ __forceinline__ void operator()() const {} };

// out1 != nullptr: the low byte of every record's code is also written to out1[offset of the record] (pass 2: digit 1
// of the record as one byte, so that pass 3 counts its bucket from 1 byte per record instead of re-reading 8)
// REV: the tiles of a range are walked from the LAST to the first: on entry sm.cursor[b] is the global offset just PAST
// bin b's records of this tile and on exit it has moved down by the tile's count (pass 3, see there).  The order inside
// a tile is unchanged, so the partition stays stable.
template <bool OUT4, int NBITS, bool FULL, typename RecFn, typename DigitFn, typename HookFn = M3NoHook, bool REV = false>
__device__ __forceinline__ void m3_partition_tile2(M3Smem2 &sm, RecFn rec, const uint32_t (&dg)[M3_ITEMS], const int BINS,
                                                   DigitFn digit, uint64_t *__restrict__ out8,
                                                   uint32_t *__restrict__ out4, HookFn after_rank = HookFn(),
                                                   uint8_t *__restrict__ out1 = nullptr) {
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t pk[M3_ITEMS];      // slot in wcnt (low 16 bits) | rank inside the (warp, bin) run (high 16); ~0u = no record
#pragma unroll
    for (int j = 0; j < M3_ITEMS; ++j) {
        const uint32_t d = dg[j];
        uint32_t same, other = 0u;
        bool ok = true;
        if (FULL) {
            same = 0xffffffffu;
        } else {
            ok = d != (uint32_t)BINS;
            same = __ballot_sync(0xffffffffu, ok);
        }
#pragma unroll
        for (int b = 0; b < NBITS; ++b) m3_vote_bit(d, 1u << b, same, other);
        const uint32_t peers = same & ~other, below = peers & lt;
        const uint32_t slot = w * M3_MAX_BINS + (FULL ? d : (d & (M3_MAX_BINS - 1)));
        const uint32_t cnt = sm.wcnt[slot];
        pk[j] = (FULL || ok) ? (slot | ((cnt + __popc(below)) << 16)) : 0xffffffffu;
        __syncwarp();
        if ((FULL || ok) && below == 0u) sm.wcnt[slot] = cnt + __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    after_rank();
    // bin tid (tid < BINS): the 8 per-warp counts -> exclusive bases; block scan over the bins
    // (the 8 counts are NOT kept in registers across the barrier below: with 16 record registers and 8 packed ranks
    // live that spilled 92-112 bytes per thread at 48 registers; they are read again when the bases are written)
    uint32_t tot = 0;
    if (tid < BINS) {
#pragma unroll
        for (int ww = 0; ww < M3_WARPS; ++ww) tot += sm.wcnt[ww * M3_MAX_BINS + tid];
    }
    uint32_t inc = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) sm.wsum[w] = inc;
    __syncthreads();
    uint32_t all = M3_TILE;
    {
        const uint4 s0 = reinterpret_cast<const uint4 *>(sm.wsum)[0], s1 = reinterpret_cast<const uint4 *>(sm.wsum)[1];
        const uint32_t v[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
        uint32_t base = 0, sum = 0;
#pragma unroll
        for (int ww = 0; ww < M3_WARPS; ++ww) {
            if (ww < w) base += v[ww];
            sum += v[ww];
        }
        if (!FULL) all = sum;
        if (tid < BINS) {
            uint32_t run = base + inc - tot;            // start of bin tid inside the tile
            const uint32_t cur = REV ? sm.cursor[tid] - tot : sm.cursor[tid];
            sm.delta[tid] = cur - run;
            sm.cursor[tid] = REV ? cur : cur + tot;
#pragma unroll
            for (int ww = 0; ww < M3_WARPS; ++ww) {
                const uint32_t cw = sm.wcnt[ww * M3_MAX_BINS + tid];
                sm.wcnt[ww * M3_MAX_BINS + tid] = run;
                run += cw;
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < M3_ITEMS; ++j)
        if (FULL || pk[j] != 0xffffffffu) sm.stage[sm.wcnt[pk[j] & 0xffffu] + (pk[j] >> 16)] = rec(j);
    __syncthreads();
    m3_clear_table2(sm);      // re-arm the table for the next tile while the stage drains
#pragma unroll
    for (int j = 0; j < M3_ITEMS; ++j) {
        const uint32_t s = j * M3_THREADS + tid;
        if (FULL || s < all) {
            const uint2 x = sm.stage[s];
            const uint32_t at = sm.delta[digit(x.y)] + s;
            if (OUT4) out4[at] = x.x; else out8[at] = ((uint64_t)x.y << 32) | x.x;
            if (!OUT4 && out1) out1[at] = (uint8_t)x.y;
        }
    }
    __syncthreads();
}

#endif

}  // namespace pdg
