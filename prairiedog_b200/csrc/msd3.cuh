// msd3.cuh -- constants and the per-tile partition core of the offsets-first MSD sort (msd3.cu).
#pragma once
#include "traverse.cuh"

namespace pdg {

static constexpr int M3_THREADS = 256;
static constexpr int M3_WARPS = M3_THREADS / 32;
static constexpr int M3_ITEMS = 16;
static constexpr int M3_TILE = M3_THREADS * M3_ITEMS;   // 4096 items per tile
static constexpr int M3_MAX_BINS = 256;
static constexpr int M3_WORDS = M3_TILE / 32 + 2;       // packed words staged per position tile
static constexpr int kMsd3MaxBits = 22;                 // k <= 11: digits (<=7, <=7, <=8)
static constexpr uint32_t M3_DROP = 0xFFFFFFFFu;        // digit of an item that is not a record

// dynamic shared memory of one tile partition
struct M3Smem {
    uint64_t stage[M3_TILE];                 // the tile in sorted order
    uint2 wtab[M3_WARPS][M3_MAX_BINS];       // per (warp, bin): {lane mask of the round, count}
    uint32_t bstart[M3_MAX_BINS];            // start of the bin inside the tile
    uint32_t delta[M3_MAX_BINS];             // global offset of the bin's run minus bstart
    uint32_t cursor[M3_MAX_BINS];            // global offset where this CTA writes the bin next
    uint32_t total;                          // records in the current tile
};

#ifdef __CUDACC__
// Stable partition of one tile by a digit.  r / dg: the items of this thread in warp-striped order
// (item j of lane l of warp w is tile element w*512 + j*32 + l); dg == M3_DROP drops the item.
// On entry sm.cursor[b] is the global offset of the first record of bin b of this tile; on exit
// it has advanced by the tile's count (so a CTA can walk consecutive tiles of one range).
// OUT4: write the low 32 bits (position) to out4, else the record to out8.
// Ranking: every lane ORs its bit into the mask of its (warp, bin) entry, reads {mask, count}
// back and the lowest lane of each group folds the mask into the count -- no warp match (MATCH.ANY
// sustains only ~0.5 lanes/cycle/SM on sm_100, shared-memory atomics ~10).
template <bool OUT4>
__device__ __forceinline__ void m3_partition_tile(M3Smem &sm, uint64_t (&r)[M3_ITEMS], uint32_t (&dg)[M3_ITEMS],
                                                  int shift, int BINS, uint64_t *__restrict__ out8,
                                                  uint32_t *__restrict__ out4) {
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    for (int i = tid; i < M3_WARPS * M3_MAX_BINS; i += M3_THREADS) (&sm.wtab[0][0])[i] = make_uint2(0u, 0u);
    __syncthreads();
    uint32_t rank[M3_ITEMS];
    uint2 *wt = sm.wtab[w];
#pragma unroll
    for (int j = 0; j < M3_ITEMS; ++j) {
        const bool ok = dg[j] != M3_DROP;
        if (ok) atomicOr(&wt[dg[j]].x, 1u << lane);
        __syncwarp();
        uint2 mc = make_uint2(0u, 0u);
        if (ok) mc = wt[dg[j]];
        rank[j] = mc.y + __popc(mc.x & lt);
        __syncwarp();
        if (ok && (mc.x & lt) == 0) wt[dg[j]] = make_uint2(0u, mc.y + __popc(mc.x));
        __syncwarp();
    }
    __syncthreads();
    // per bin: exclusive offsets of the warps, tile count
    for (int b = tid; b < BINS; b += M3_THREADS) {
        uint32_t run = 0;
#pragma unroll
        for (int ww = 0; ww < M3_WARPS; ++ww) {
            const uint32_t c = sm.wtab[ww][b].y;
            sm.wtab[ww][b].y = run;
            run += c;
        }
        sm.bstart[b] = run;
    }
    __syncthreads();
    if (w == 0) {   // exclusive scan of the bin counts by one warp
        constexpr int PER = M3_MAX_BINS / 32;
        uint32_t v[PER], s = 0;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            const int b = lane * PER + i;
            v[i] = b < BINS ? sm.bstart[b] : 0u;
            s += v[i];
        }
        uint32_t inc = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        uint32_t ex = inc - s;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            const int b = lane * PER + i;
            if (b < BINS) {
                sm.bstart[b] = ex;
                const uint32_t cur = sm.cursor[b];
                sm.delta[b] = cur - ex;
                sm.cursor[b] = cur + v[i];
                ex += v[i];
            }
        }
        if (lane == 31) sm.total = ex;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < M3_ITEMS; ++j)
        if (dg[j] != M3_DROP) sm.stage[sm.bstart[dg[j]] + wt[dg[j]].y + rank[j]] = r[j];
    __syncthreads();
    const uint32_t total = sm.total, mask = BINS - 1;
#pragma unroll
    for (int j = 0; j < M3_ITEMS; ++j) {
        const uint32_t s = j * M3_THREADS + tid;
        if (s < total) {
            const uint64_t x = sm.stage[s];
            const uint32_t d = ((uint32_t)(x >> 32) >> shift) & mask;
            if (OUT4) out4[sm.delta[d] + s] = (uint32_t)x; else out8[sm.delta[d] + s] = x;
        }
    }
    __syncthreads();
}
#endif

}  // namespace pdg
