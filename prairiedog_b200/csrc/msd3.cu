// msd3.cu -- offsets-first MSD sort of the k-mer occurrence index for k <= 11.
//
// The 2k-bit code is cut into three digits d3 | d2 | d1 (most significant first).  ONE pass over
// the 2-bit packed genome counts (a) d3 per 4096-position tile and (b) the joint (d3, d2)
// histogram per chunk of tiles.  Prefix sums of those tables give every partition step its exact
// output offsets in advance, so no pass needs look-back, tile status words or a host round trip:
//   pass 1  position tile  -> records (code << 32 | position) grouped by d3           0.25 + 8 B
//   pass 2  range (chunk, d3) of pass 1's output -> grouped by (d3, d2)               8 + 8 B
//   pass 3  bucket (d3, d2): count d1, emit the row starts of its 2^b1 codes, then    8 + 4 B
//           write the position column grouped by the full code (second read hits L2)
// Every partition is stable, so positions ascend inside a row.  Row starts for ALL 4^k codes come
// out of pass 3 (the exclusive scan of the full code histogram), no per-record atomics needed.
//
// Reference counterpart: lemongraph/lib/lemongraph.c:591-630 (node_idx / srcnode_idx /
// tgtnode_idx are kept ordered by one B-tree insert per key).
#include "gindex.cuh"
#include "msd3.cuh"
#include <algorithm>
#include <stdlib.h>

namespace pdg {

struct M3Plan {
    int b1, b2, b3;              // digit widths, low to high
    uint32_t n_tiles, ch, n_chunks;
    uint32_t n_chunks_cap;       // no sub-range of these positions plans more chunks than this (buffer sizing)
    uint64_t base0;              // first position of tile 0 (a multiple of 32)
};

// ---- per-tile contig lookup ----------------------------------------------------------------------
// tile_info[tile] = index of the last contig that begins at or before the tile (0 if none), with
// M3_SIMPLE set when every window of the tile is a k-mer of that contig.
__global__ void m3_tile_info_kernel(ContigTable ct, M3Plan pl, int kv, uint32_t *__restrict__ tile_info) {
    const uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile >= pl.n_tiles) return;
    const uint64_t p0 = pl.base0 + (uint64_t)tile * M3_TILE;
    const uint64_t ci = contig_of(ct, p0);
    const bool simple = p0 >= ct.begin[ci] && p0 + M3_TILE - 1 + kv <= ct.end[ci];
    tile_info[tile] = (uint32_t)ci | (simple ? M3_SIMPLE : 0u);
}

// ---- histograms from the packed genome --------------------------------------------------------
// H1T[d3][tile]; H12[chunk][d3 << b2 | d2]; tot32[cell] += chunk counts.
// One CTA per chunk; every WARP takes whole tiles on its own (private d3 counters), so there is no block barrier
// inside the tile loop.  A LANE owns one packed word (32 consecutive positions) and walks it from registers: the
// window at position i of the word is one funnel shift by a compile-time amount of (word, top half of the next
// word), its two digits one shift each.  (Round 1 walked the positions lane-parallel instead -- lane l = position
// 32 i + l -- which cost three shuffles per round to broadcast word i: 30 instructions per 32 positions, 0.65 ms.)
static constexpr int M3H_THREADS = 1024;
static constexpr int M3H_WARPS = M3H_THREADS / 32;
__global__ void __launch_bounds__(M3H_THREADS) m3_hist_kernel(const uint2 *__restrict__ packed, uint64_t packed_words,
                                                              int k, int kv, ContigTable ct, M3Plan pl,
                                                              const uint32_t *__restrict__ tile_info,
                                                              uint32_t *__restrict__ H1T, uint32_t *__restrict__ H12,
                                                              unsigned int *__restrict__ tot32) {
    extern __shared__ uint32_t h12[];   // [cells]
    __shared__ uint32_t h1[M3H_WARPS][128];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t cells = 1u << (pl.b3 + pl.b2), B3 = 1u << pl.b3;
    // a window is the top 2k bits of a 32-bit funnel: its digits are plain right shifts of the funnel
    const int sh2 = 32 - 2 * k + pl.b1, sh3 = 32 - 2 * k + pl.b1 + pl.b2;
    for (uint32_t i = tid; i < cells; i += M3H_THREADS) h12[i] = 0;
    __syncthreads();
    const uint32_t t0 = blockIdx.x * pl.ch, t1 = min(t0 + pl.ch, pl.n_tiles);
    for (uint32_t tile = t0 + w; tile < t1; tile += M3H_WARPS) {
        const uint64_t w0 = (pl.base0 >> 5) + (uint64_t)tile * M3_TILE_WORDS;
        const uint32_t info = tile_info[tile];
        for (int i = lane; i < 128; i += 32) h1[w][i] = 0;
        const bool simple = (info & M3_SIMPLE) != 0;
        __syncwarp();
#pragma unroll 1
        for (int g = 0; g < M3_TILE_WORDS / 32; ++g) {
            const uint64_t wa = w0 + 32 * g + lane;
            const uint2 mine = wa < packed_words ? packed[wa] : make_uint2(0u, 0u);
            // only the top half of the following word is ever needed (k <= 11 <= 16 residues)
            const uint32_t nxt = wa + 1 < packed_words ? packed[wa + 1].y : 0u;
            if (simple) {
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const uint32_t f = i < 16 ? __funnelshift_l(mine.x, mine.y, 2 * i) : __funnelshift_l(nxt, mine.x, 2 * (i - 16));
                    atomicAdd(&h1[w][f >> sh3], 1u);
                    atomicAdd(&h12[f >> sh2], 1u);
                }
            } else {
                const uint32_t vmask = m3_valid_word(ct, wa, info & ~M3_SIMPLE, kv);
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const uint32_t f = i < 16 ? __funnelshift_l(mine.x, mine.y, 2 * i) : __funnelshift_l(nxt, mine.x, 2 * (i - 16));
                    if ((vmask >> i) & 1u) {
                        atomicAdd(&h1[w][f >> sh3], 1u);
                        atomicAdd(&h12[f >> sh2], 1u);
                    }
                }
            }
        }
        __syncwarp();
        for (uint32_t i = lane; i < B3; i += 32) H1T[(size_t)i * pl.n_tiles + tile] = h1[w][i];
        __syncwarp();
    }
    __syncthreads();
    for (uint32_t i = tid; i < cells; i += M3H_THREADS) {
        const uint32_t v = h12[i];
        H12[(size_t)blockIdx.x * cells + i] = v;
        if (v) atomicAdd(&tot32[i], v);
    }
}

// base32[cell] = exclusive scan of tot32 (cells + 1 entries; the last is n).  One CTA.
__global__ void __launch_bounds__(1024) m3_base_kernel(const unsigned int *__restrict__ tot32, uint32_t cells,
                                                       uint32_t *__restrict__ base32, uint32_t *__restrict__ note,
                                                       uint32_t note_value) {
    if (note && threadIdx.x == 0) *note = note_value;   // a group's record offset, for pass 3
    uint64_t carry = 0;
    for (uint32_t b = 0; b < cells; b += 1024) {
        const uint32_t i = b + threadIdx.x;
        const uint64_t v = i < cells ? tot32[i] : 0;
        uint64_t tot;
        const uint64_t ex = block_excl_scan(v, &tot);
        if (i < cells) base32[i] = (uint32_t)(carry + ex);
        carry += tot;
    }
    if (threadIdx.x == 0) base32[cells] = (uint32_t)carry;
}

// H1T[d3][tile] -> global offset of the first record of (tile, d3) in pass 1's output, in place.
// One CTA per chunk, a warp per d3.  The range (chunk, d3) starts at
//   base32[d3 << b2] + sum over d2 of (O2[chunk][d3, d2] - base32[d3, d2])
// (O2 = H12 after m3_scan_chunks_kernel: records of cell (d3, d2) in earlier chunks), so every
// (chunk, d3) pair scans its own tiles independently.
__global__ void __launch_bounds__(256) m3_scan_tiles_kernel(uint32_t *__restrict__ H1T, M3Plan pl,
                                                            const uint32_t *__restrict__ O2,
                                                            const uint32_t *__restrict__ base32) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t B2 = 1u << pl.b2, B3 = 1u << pl.b3, c = blockIdx.x;
    const uint32_t t0 = c * pl.ch, t1 = min(t0 + pl.ch, pl.n_tiles);
    for (uint32_t d3 = w; d3 < B3; d3 += 8) {
        uint32_t acc = 0;
        for (uint32_t d2 = lane; d2 < B2; d2 += 32) {
            const uint32_t cell = (d3 << pl.b2) | d2;
            acc += O2[(size_t)c * (B3 * B2) + cell] - base32[cell];
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        uint32_t carry = base32[(size_t)d3 << pl.b2] + acc;
        uint32_t *col = H1T + (size_t)d3 * pl.n_tiles;
        for (uint32_t t = t0; t < t1; t += 32) {
            const uint32_t i = t + lane;
            const uint32_t v = i < t1 ? col[i] : 0u;
            uint32_t inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t x = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += x;
            }
            if (i < t1) col[i] = carry + inc - v;
            carry += __shfl_sync(0xffffffffu, inc, 31);
        }
    }
}

// per cell: H12[chunk][cell] -> base32[cell] + exclusive prefix over chunks, in place
__global__ void __launch_bounds__(256) m3_scan_chunks_kernel(uint32_t *__restrict__ H12, uint32_t n_chunks, uint32_t cells,
                                                             const uint32_t *__restrict__ base32) {
    const uint32_t cell = blockIdx.x * 256 + threadIdx.x;
    if (cell >= cells) return;
    uint32_t run = base32[cell];
    constexpr int U = 16;   // independent loads in flight
    for (uint32_t c0 = 0; c0 < n_chunks; c0 += U) {
        uint32_t v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = c0 + u < n_chunks ? H12[(size_t)(c0 + u) * cells + cell] : 0u;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (c0 + u < n_chunks) H12[(size_t)(c0 + u) * cells + cell] = run;
            run += v[u];
        }
    }
}

// ---- pass 1: position tile -> records grouped by d3 ---------------------------------------------
// FULL tile = every window of the tile is a k-mer (tile_info's M3_SIMPLE): no validity words, no guards.
// A CTA walks M3_P1_TILES consecutive tiles: in pass 1's output the records of a digit are ordered by position, so
// the run of (tile + 1, d3) starts where the run of (tile, d3) ends and the cursor the core advances is already right
// -- the 128 scattered words of O1T (one per digit, n_tiles words apart) are fetched once per CTA, not once per tile.
#ifndef M3_P1_TILES
#define M3_P1_TILES 1   /* measured on B200: 1 tile 2.044 ms, 2: 2.096, 4: 2.069, 8: 2.075 -- the cursor loads were hidden anyway */
#endif
template <int NBITS>
__global__ void __launch_bounds__(M3_THREADS, M3_MINB) m3_pass1_kernel(const uint2 *__restrict__ packed,
                                                                 uint64_t packed_words, int k, int kv, ContigTable ct, M3Plan pl,
                                                                 const uint32_t *__restrict__ tile_info,
                                                                 const uint32_t *__restrict__ O1T,
                                                                 uint64_t *__restrict__ out) {
    extern __shared__ __align__(16) unsigned char m3_raw[];
    M3Smem2 &sm = *reinterpret_cast<M3Smem2 *>(m3_raw);
    __shared__ uint2 words[M3_TILE_WORDS + 2];
    __shared__ uint32_t vm[M3_TILE_WORDS];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t tile0 = blockIdx.x * M3_P1_TILES, tile1 = min(tile0 + (uint32_t)M3_P1_TILES, pl.n_tiles);
    const uint32_t B3 = 1u << pl.b3;
    const int sh3 = pl.b1 + pl.b2;
    for (uint32_t b = tid; b < B3; b += M3_THREADS) sm.cursor[b] = O1T[(size_t)b * pl.n_tiles + tile0];
    m3_clear_table2(sm);
    const uint2 *wrow = words + w * M3_ITEMS;
    auto dig = [sh3](uint32_t c) { return c >> sh3; };
    static_assert(M3_TILE_WORDS <= M3_THREADS, "one validity word per thread");
    for (uint32_t tile = tile0; tile < tile1; ++tile) {
        const uint64_t w0 = (pl.base0 >> 5) + (uint64_t)tile * M3_TILE_WORDS;
        const uint32_t info = tile_info[tile];
        const bool full = (info & M3_SIMPLE) != 0;          // the same for the whole CTA
        // (the previous tile's partition ended with a barrier: nobody reads words / vm any more)
        for (int i = tid; i < M3_TILE_WORDS + 2; i += M3_THREADS)
            words[i] = w0 + i < packed_words ? packed[w0 + i] : make_uint2(0u, 0u);
        if (!full && tid < M3_TILE_WORDS) vm[tid] = m3_valid_word(ct, w0 + tid, info & ~M3_SIMPLE, kv);
        __syncthreads();
        const uint32_t pbase = (uint32_t)(w0 << 5) + w * M3_WARP_TILE + lane;
        uint32_t dg[M3_ITEMS];
#ifdef M3_P1_RECOMPUTE
        // round 1: only the digit stays in registers, the code is formed again from the staged words at scatter time
        auto rec = [&](int j) { return make_uint2(pbase + 32 * j, m3_code_at(wrow[j], wrow[j + 1], lane, k)); };
        if (full) {
#pragma unroll
            for (int j = 0; j < M3_ITEMS; ++j) dg[j] = m3_code_at(wrow[j], wrow[j + 1], lane, k) >> sh3;
            m3_partition_tile2<false, NBITS, true>(sm, rec, dg, (int)B3, dig, out, nullptr);
        } else {
#pragma unroll
            for (int j = 0; j < M3_ITEMS; ++j)
                dg[j] = ((vm[w * M3_ITEMS + j] >> lane) & 1u) ? (m3_code_at(wrow[j], wrow[j + 1], lane, k) >> sh3) : B3;
            m3_partition_tile2<false, NBITS, false>(sm, rec, dg, (int)B3, dig, out, nullptr);
        }
#else
        // the 8 codes stay in registers (pass 1 holds no records, so there is room at 48 registers): forming them
        // again at scatter time cost two broadcast LDS.64 per item, and ncu shows pass 1 stalled on shared-memory
        // latency (short_scoreboard 3.5 per issue, profiles/r02_ncu_full_k11_summary.txt)
        uint32_t code[M3_ITEMS];
#pragma unroll
        for (int j = 0; j < M3_ITEMS; ++j) code[j] = m3_code_at(wrow[j], wrow[j + 1], lane, k);
        auto rec = [&](int j) { return make_uint2(pbase + 32 * j, code[j]); };
        if (full) {
#pragma unroll
            for (int j = 0; j < M3_ITEMS; ++j) dg[j] = code[j] >> sh3;
            m3_partition_tile2<false, NBITS, true>(sm, rec, dg, (int)B3, dig, out, nullptr);
        } else {
#pragma unroll
            for (int j = 0; j < M3_ITEMS; ++j) dg[j] = ((vm[w * M3_ITEMS + j] >> lane) & 1u) ? (code[j] >> sh3) : B3;
            m3_partition_tile2<false, NBITS, false>(sm, rec, dg, (int)B3, dig, out, nullptr);
        }
#endif
    }
}

// ---- pass 2: range (chunk, d3) -> grouped by (d3, d2) --------------------------------------------
// The tiles of a range are fetched by bulk asynchronous copies (tma.cuh) into `inbuf`, one tile ahead: the copy of
// tile t + 1 is started by one thread as soon as every thread holds tile t in registers (after the ranking barrier)
// and lands while tile t is scanned, scattered and written.
template <int NBITS>
__global__ void __launch_bounds__(M3_THREADS, M3_MINB) m3_pass2_kernel(const uint2 *__restrict__ in, M3Plan pl,
                                                                 const uint32_t *__restrict__ O1T,
                                                                 const uint32_t *__restrict__ O2,
                                                                 const uint32_t *__restrict__ base32,
                                                                 uint64_t *__restrict__ out, uint8_t *__restrict__ out_d1) {
    extern __shared__ __align__(16) unsigned char m3_raw[];
    M3Smem2 &sm = *reinterpret_cast<M3Smem2 *>(m3_raw);
    __shared__ __align__(16) uint2 inbuf[M3_TILE + 2];
    __shared__ __align__(8) unsigned long long bar;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t B2 = 1u << pl.b2, B3 = 1u << pl.b3;
    const uint32_t c = blockIdx.x / B3, d3 = blockIdx.x % B3;
    const uint32_t begin = O1T[(size_t)d3 * pl.n_tiles + (size_t)c * pl.ch];
    const uint32_t end = (c + 1 < pl.n_chunks) ? O1T[(size_t)d3 * pl.n_tiles + (size_t)(c + 1) * pl.ch]
                                                : base32[(size_t)(d3 + 1) << pl.b2];
    if (begin >= end) return;
    if (tid == 0) {
        tma_bar_init(&bar, 1);
        const TmaSpan sp = tma_span8(in + begin, min((uint32_t)M3_TILE, end - begin));
        tma_load(inbuf, sp.src, sp.bytes, &bar);
    }
    for (uint32_t b = tid; b < B2; b += M3_THREADS) sm.cursor[b] = O2[(size_t)c * (B3 * B2) + ((size_t)d3 << pl.b2) + b];
    m3_clear_table2(sm);
    __syncthreads();
    const int b1 = pl.b1;
    const uint32_t m2 = B2 - 1;
    const uint32_t li0 = w * M3_WARP_TILE + lane;
    auto dig = [b1, m2](uint32_t cd) { return (cd >> b1) & m2; };
    uint32_t parity = 0;
    for (uint32_t t0 = begin; t0 < end; t0 += M3_TILE) {
        const uint32_t tile_n = min((uint32_t)M3_TILE, end - t0);
        const uint2 *src = inbuf + tma_span8(in + t0, tile_n).skip + li0;
        // the next tile of the range, fetched while this one is processed
        const uint32_t nt0 = t0 + M3_TILE;
        auto prefetch = [&]() {
            if (tid == 0 && nt0 < end) {
                const TmaSpan sp = tma_span8(in + nt0, min((uint32_t)M3_TILE, end - nt0));
                tma_load(inbuf, sp.src, sp.bytes, &bar);
            }
        };
        uint2 r[M3_ITEMS];
        uint32_t dg[M3_ITEMS];
        tma_bar_wait(&bar, parity);
        parity ^= 1u;
        if (tile_n == (uint32_t)M3_TILE) {          // full tile: straight-line loads, no sentinel
#pragma unroll
            for (int j = 0; j < M3_ITEMS; ++j) r[j] = src[32 * j];
#pragma unroll
            for (int j = 0; j < M3_ITEMS; ++j) dg[j] = (r[j].y >> b1) & m2;
            auto rec = [&](int j) { return r[j]; };
            m3_partition_tile2<false, NBITS, true>(sm, rec, dg, (int)B2, dig, out, nullptr, prefetch, out_d1);
        } else {
#pragma unroll
            for (int j = 0; j < M3_ITEMS; ++j) {
                const bool ok = li0 + 32 * j < tile_n;
                r[j] = ok ? src[32 * j] : make_uint2(0u, 0u);
                dg[j] = ok ? ((r[j].y >> b1) & m2) : B2;
            }
            auto rec = [&](int j) { return r[j]; };
            m3_partition_tile2<false, NBITS, false>(sm, rec, dg, (int)B2, dig, out, nullptr, prefetch, out_d1);
        }
    }
}

// ---- pass 3: bucket (d3, d2): count d1, row starts, position column ------------------------------
// The bucket of a cell is the concatenation, in group order, of the cell's segment of every group
// (a group = one run of passes 1-2 over a range of positions, see msd3_stage_group):
// segment g = in[goff[g] + gbase[g][cell] .. goff[g] + gbase[g][cell + 1]).
// tot[cell] = records of the cell over all groups (input of the base scan)
__global__ void m3_cell_totals_kernel(const uint32_t *__restrict__ gbase, uint32_t n_groups, uint32_t cells,
                                      unsigned int *__restrict__ tot) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= cells) return;
    uint32_t t = 0;
    for (uint32_t g = 0; g < n_groups; ++g) {
        const uint32_t *gb = gbase + (size_t)g * (cells + 1);
        t += gb[c + 1] - gb[c];
    }
    tot[c] = t;
}

#ifndef M3_COUNT_UNROLL
#define M3_COUNT_UNROLL 8   /* loads in flight per thread in the count loop; 1, 2, 4: 2-8% slower on B200 */
#endif
template <int NBITS>
__global__ void __launch_bounds__(M3_THREADS, M3_MINB3) m3_pass3_kernel(const uint2 *__restrict__ in, M3Plan pl,
                                                                 const uint32_t *__restrict__ gbase,
                                                                 const uint32_t *__restrict__ goff, uint32_t n_groups,
                                                                 const uint32_t *__restrict__ base32,
                                                                 uint32_t *__restrict__ rowtab,
                                                                 uint32_t *__restrict__ out_pos,
                                                                 const uint8_t *__restrict__ in_d1) {
    extern __shared__ __align__(16) unsigned char m3_raw[];
    M3Smem2 &sm = *reinterpret_cast<M3Smem2 *>(m3_raw);
    __shared__ __align__(16) uint2 inbuf[M3_TILE + 2];
    __shared__ __align__(8) unsigned long long bar;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t B1 = 1u << pl.b1, m1 = B1 - 1, cell = blockIdx.x;
    const uint32_t cells1 = (1u << (pl.b3 + pl.b2)) + 1;
    const uint32_t begin = base32[cell];
    // count d1 over the bucket (sm.delta doubles as the histogram)
    for (int b = tid; b < M3_MAX_BINS; b += M3_THREADS) sm.delta[b] = 0;
    m3_clear_table2(sm);
    __syncthreads();
    for (uint32_t gi = 0; gi < n_groups; ++gi) {
        const uint32_t *gb = gbase + (size_t)gi * cells1;
        const uint32_t sb = goff[gi] + gb[cell], se = goff[gi] + gb[cell + 1];
        if (in_d1) {
            // Counted from pass 2's side array: ONE byte per record (the low byte of its code).  Counting from the
            // records reads the bucket twice -- once here, once for the partition below -- and at config-3 scale the
            // second read no longer hits L2 (592 buckets of 244 KB in flight > 126 MB): ncu measured 21.2 B of DRAM
            // traffic per record against 12 algorithmic.  The side array costs pass 2 5e8 scattered byte stores
            // (2.14 -> 2.90 ms), so the host uses it only where pass 2 is hidden and pass 3 exposed: the pipelined
            // build (profiles/r02_tune_d1byte.txt).
            // Word-wise: thread t takes the aligned 32-bit words t, t + 256, ... of the segment; words at the two
            // ends are masked by index.
            const uint32_t w_first = sb >> 2, w_last = (se + 3) >> 2;          // words [w_first, w_last)
            const uint32_t *d32 = reinterpret_cast<const uint32_t *>(in_d1);
            uint32_t wd = w_first + tid;
            for (; wd + (M3_COUNT_UNROLL - 1) * M3_THREADS < w_last; wd += M3_COUNT_UNROLL * M3_THREADS) {
                uint32_t v[M3_COUNT_UNROLL];
#pragma unroll
                for (int u = 0; u < M3_COUNT_UNROLL; ++u) v[u] = d32[wd + u * M3_THREADS];
#pragma unroll
                for (int u = 0; u < M3_COUNT_UNROLL; ++u) {
                    const uint32_t i0 = (wd + u * M3_THREADS) << 2;
                    if (i0 >= sb && i0 + 3 < se) {
#pragma unroll
                        for (int b = 0; b < 4; ++b) atomicAdd(&sm.delta[(v[u] >> (8 * b)) & m1], 1u);
                    } else {
#pragma unroll
                        for (int b = 0; b < 4; ++b)
                            if (i0 + b >= sb && i0 + b < se) atomicAdd(&sm.delta[(v[u] >> (8 * b)) & m1], 1u);
                    }
                }
            }
            for (; wd < w_last; wd += M3_THREADS) {
                const uint32_t v = d32[wd], i0 = wd << 2;
#pragma unroll
                for (int b = 0; b < 4; ++b)
                    if (i0 + b >= sb && i0 + b < se) atomicAdd(&sm.delta[(v >> (8 * b)) & m1], 1u);
            }
        } else {
            // from the records themselves: whole blocks of M3_COUNT_UNROLL x M3_THREADS records without a guard, the
            // tail one record per thread and trip
            uint32_t i = sb + tid;
            for (; i + (M3_COUNT_UNROLL - 1) * M3_THREADS < se; i += M3_COUNT_UNROLL * M3_THREADS) {
                uint32_t cd[M3_COUNT_UNROLL];
#pragma unroll
                for (int u = 0; u < M3_COUNT_UNROLL; ++u) cd[u] = in[i + u * M3_THREADS].y;
#pragma unroll
                for (int u = 0; u < M3_COUNT_UNROLL; ++u) atomicAdd(&sm.delta[cd[u] & m1], 1u);
            }
            for (; i < se; i += M3_THREADS) atomicAdd(&sm.delta[in[i].y & m1], 1u);
        }
    }
    __syncthreads();
    {   // exclusive scan of the 256 counts -> first occurrence of every code of the bucket
        uint32_t v[M3_BPT], sum = 0;
#pragma unroll
        for (int i = 0; i < M3_BPT; ++i) {
            const int b = tid * M3_BPT + i;
            v[i] = b < M3_MAX_BINS ? sm.delta[b] : 0u;
            sum += v[i];
        }
        uint32_t inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) sm.wsum[w] = inc;
        __syncthreads();
        uint32_t ex = begin + inc - sum;
#pragma unroll
        for (int ww = 0; ww < M3_WARPS; ++ww)
            if (ww < w) ex += sm.wsum[ww];
#pragma unroll
        for (int i = 0; i < M3_BPT; ++i) {
            const int b = tid * M3_BPT + i;
            if (b < (int)B1) rowtab[((size_t)cell << pl.b1) + b] = ex;
            ex += v[i];
            if (b < M3_MAX_BINS) sm.cursor[b] = ex;     // the END of the row: the tiles are walked backwards (below)
        }
    }
    __syncthreads();
    const uint32_t li0 = w * M3_WARP_TILE + lane;
    auto dig = [m1](uint32_t cd) { return cd & m1; };
    // The tiles of the bucket are walked BACKWARDS -- last segment first, inside a segment from its end -- with cursors
    // that start at the end of every row and move down.  The count sweep above read the bucket front to back; walking
    // it front to back again is the worst case for the L2: 592 buckets of 244 KB are in flight against 126 MB, so
    // by the time the second sweep starts the front of the bucket is the oldest data and gone (ncu: 21.2 B of DRAM
    // traffic per record against 12 algorithmic).  Backwards, the second sweep starts on what was read last.
    // Tiles are fetched one ahead by bulk asynchronous copies (see pass 2); every thread walks the same sequence,
    // thread 0 issues the copies.  (gi, sb, hi) = group of the current tile, start of its segment, end of the tile.
    auto segment = [&](uint32_t g, uint32_t &b0, uint32_t &e0) {
        const uint32_t *gb = gbase + (size_t)g * cells1;
        b0 = goff[g] + gb[cell];
        e0 = goff[g] + gb[cell + 1];
    };
    // the tile that ends at or before (g, h) -- skips empty segments; false when the bucket is exhausted
    auto settle = [&](int &g, uint32_t &b0, uint32_t &h) {
        while (g >= 0) {
            if (h > b0) return true;
            if (--g >= 0) segment((uint32_t)g, b0, h);
        }
        return false;
    };
    int gi = (int)n_groups - 1;
    uint32_t sb = 0, hi = 0;
    if (gi >= 0) segment((uint32_t)gi, sb, hi);
    bool have = settle(gi, sb, hi);
    if (tid == 0) {
        tma_bar_init(&bar, 1);
        if (have) {
            const uint32_t n0 = min((uint32_t)M3_TILE, hi - sb);
            const TmaSpan sp = tma_span8(in + (hi - n0), n0);
            tma_load(inbuf, sp.src, sp.bytes, &bar);
        }
    }
    __syncthreads();
    uint32_t parity = 0;
    while (have) {
        const uint32_t tile_n = min((uint32_t)M3_TILE, hi - sb), t0 = hi - tile_n;
        const uint2 *src = inbuf + tma_span8(in + t0, tile_n).skip + li0;
        int ngi = gi;
        uint32_t nsb = sb, nhi = t0;
        const bool nhave = settle(ngi, nsb, nhi);
        auto prefetch = [&]() {
            if (tid == 0 && nhave) {
                const uint32_t n1 = min((uint32_t)M3_TILE, nhi - nsb);
                const TmaSpan sp = tma_span8(in + (nhi - n1), n1);
                tma_load(inbuf, sp.src, sp.bytes, &bar);
            }
        };
        uint2 r[M3_ITEMS];
        uint32_t dg[M3_ITEMS];
        tma_bar_wait(&bar, parity);
        parity ^= 1u;
        if (tile_n == (uint32_t)M3_TILE) {
#pragma unroll
            for (int j = 0; j < M3_ITEMS; ++j) r[j] = src[32 * j];
#pragma unroll
            for (int j = 0; j < M3_ITEMS; ++j) dg[j] = r[j].y & m1;
            auto rec = [&](int j) { return r[j]; };
            m3_partition_tile2<true, NBITS, true, decltype(rec), decltype(dig), decltype(prefetch), true>(
                sm, rec, dg, (int)B1, dig, nullptr, out_pos, prefetch);
        } else {
#pragma unroll
            for (int j = 0; j < M3_ITEMS; ++j) {
                const bool ok = li0 + 32 * j < tile_n;
                r[j] = ok ? src[32 * j] : make_uint2(0u, 0u);
                dg[j] = ok ? (r[j].y & m1) : B1;
            }
            auto rec = [&](int j) { return r[j]; };
            m3_partition_tile2<true, NBITS, false, decltype(rec), decltype(dig), decltype(prefetch), true>(
                sm, rec, dg, (int)B1, dig, nullptr, out_pos, prefetch);
        }
        gi = ngi; sb = nsb; hi = nhi; have = nhave;
    }
}

// ---- rows from the table of row starts -----------------------------------------------------------
__global__ void m3_present_kernel(const uint32_t *__restrict__ rowtab, uint64_t n_codes, uint32_t *__restrict__ flags) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_codes) flags[i] = rowtab[i + 1] > rowtab[i] ? 1u : 0u;
}
__global__ void m3_compact_rows_kernel(const uint32_t *__restrict__ rowtab, const uint32_t *__restrict__ flags,
                                       const uint32_t *__restrict__ rank, uint64_t n_codes,
                                       uint64_t *__restrict__ codes, uint32_t *__restrict__ rowptr) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_codes && flags[i]) {
        codes[rank[i]] = i;
        rowptr[rank[i]] = rowtab[i];
    }
    if (i == 0) rowptr[rank[n_codes]] = rowtab[n_codes];
}

// ---- host -----------------------------------------------------------------------------------------
// The build runs in three steps so that FASTA batches can be staged while later files are still
// being copied (ingest_batch.cu):
//   msd3_reserve      buffers for positions [pos_lo, pos_hi) and up to max_groups groups
//   msd3_stage_group  histogram + scans + passes 1-2 over one range of positions (a group); its
//                     records land in recB at rec_off, grouped by (d3, d2); its cell bases are kept
//   msd3_finish       pass 3 over the cells (segments of all groups in order) + rows
static M3Plan m3_plan(int k, uint64_t pos0, uint64_t pos1) {
    M3Plan pl{};
    pl.b1 = std::min(M3_B1, 2 * k);
    pl.b2 = std::min(M3_B2, 2 * k - pl.b1);
    pl.b3 = 2 * k - pl.b1 - pl.b2;
    pl.base0 = pos0 & ~31ULL;
    pl.n_tiles = (uint32_t)((pos1 - pl.base0 + M3_TILE - 1) / M3_TILE);
    // ranges of pass 2 average ~m3_range records (a few tiles each: little padding, still >> SMs ranges)
    static const uint32_t m3_range = getenv("PDG_M3_RANGE") ? (uint32_t)atoi(getenv("PDG_M3_RANGE")) : 6000u;
    pl.ch = std::max<uint32_t>(1, (uint32_t)(((uint64_t)m3_range << pl.b3) / M3_TILE));
    // a small range of positions (one group of a pipelined ingest) still has to fill the GPU: the histogram
    // kernel runs one CTA per chunk
    static const uint32_t m3_min_chunks = getenv("PDG_M3_MIN_CHUNKS") ? (uint32_t)atoi(getenv("PDG_M3_MIN_CHUNKS")) : 148u;
    pl.ch = std::max<uint32_t>(1, std::min<uint32_t>(pl.ch, (pl.n_tiles + m3_min_chunks - 1) / m3_min_chunks));
    pl.n_chunks = (pl.n_tiles + pl.ch - 1) / pl.ch;
    // (Rounding the chunk count to whole waves of the histogram kernel -- 651 -> 592 chunks at config 3 -- was tried:
    // pass-2 ranges grew from 2.9 to 3.2 tiles, i.e. a nearly empty fourth tile each, 2.32 -> 2.47 ms.  The fit of a
    // range to whole tiles matters more than the histogram's last wave; profiles/r02_tune_core_v2_v3.txt.)
    pl.n_chunks_cap = pl.n_chunks + 1;
    return pl;
}

void msd3_reserve(pdg_graph *g, uint64_t pos_lo, uint64_t pos_hi, uint32_t max_groups) {
    cudaStream_t s = g->stream;
    const int kv = g->k, k = std::min(g->k, kMsd3MaxBits / 2);   // sort key: the first k residues of a kv-mer
    (void)kv;
    const M3Plan pl = m3_plan(k, pos_lo, pos_hi);
    const uint32_t B3 = 1u << pl.b3, cells = 1u << (pl.b3 + pl.b2);
    const uint64_t n_codes = 1ULL << (2 * k), n = pos_hi - pos_lo;
    // a group never has more tiles / chunks than the whole range (+1 for its own alignment)
    g->status.reserve((size_t)B3 * (pl.n_tiles + 1) * 4, s);                                          // H1T -> O1T
    g->scratch1.reserve(std::max<size_t>((size_t)(pl.n_chunks_cap + 1) * cells * 4, (n_codes + 1) * 4), s);   // H12 -> O2, later ranks
    g->hist.reserve((size_t)(2 * cells + 2) * 4, s);                                                  // tot32 | base32
    g->scratch0.reserve((n_codes + 1) * 4, s);                                                        // row starts of every code
    // + 64: a bulk copy (tma.cuh) rounds its span to 16 bytes, i.e. reads up to one record past the last one
    g->recA.reserve(std::max<size_t>(n * 8, n_codes * 4) + 64, s);
    g->recB.reserve(std::max<size_t>(n * 8, 16) + 64, s);
    // digit 1 of every record of recB as one byte, for pass 3's count: only when the build is pipelined (several
    // groups: pass 2 then runs under the host->device copies and pass 3 is the exposed tail; see m3_pass3_kernel)
    g->m3_use_d1 = max_groups > 1;
    if (g->m3_use_d1) g->m3_d1.reserve(n + 64, s);
    g->g_pos.reserve(std::max<size_t>(n * 4, 16), s);
    g->m3_tile_info.reserve((size_t)(pl.n_tiles + 1) * 4, s);
    g->m3_gbase.reserve((size_t)max_groups * (cells + 1) * 4 + (size_t)max_groups * 4, s);
    g->m3_groups.clear();
    g->m3_launched = false;
    g->m3_max_groups = max_groups;
    g->m3_pos_lo = pos_lo;
    if (!(g->attr_mask & pdg_graph::ATTR_MSD3)) {
        PDG_CUDA(cudaFuncSetAttribute(m3_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 << (22 - M3_B1)));
        PDG_CUDA(cudaFuncSetAttribute(m3_pass1_kernel<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(M3Smem2)));
        PDG_CUDA(cudaFuncSetAttribute(m3_pass2_kernel<M3_B2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(M3Smem2)));
        PDG_CUDA(cudaFuncSetAttribute(m3_pass3_kernel<M3_B1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(M3Smem2)));
        g->attr_mask |= pdg_graph::ATTR_MSD3;
    }
}

void msd3_stage_group(pdg_graph *g, const ContigTable &ct, uint64_t pos0, uint64_t pos1) {
    cudaStream_t s = g->stream;
    const int kv = g->k, k = std::min(g->k, kMsd3MaxBits / 2);   // sort key: the first k residues of a kv-mer
    (void)kv;
    if (g->m3_groups.size() >= g->m3_max_groups) {
        set_error("msd3_stage_group: more than %u groups", g->m3_max_groups);
        throw CudaError{PDG_E_INVALID};
    }
    const M3Plan pl = m3_plan(k, pos0, pos1);
    const uint32_t B3 = 1u << pl.b3, cells = 1u << (pl.b3 + pl.b2);
    const uint32_t gi = (uint32_t)g->m3_groups.size();
    const uint64_t rec_off = pos0 - g->m3_pos_lo;     // a group has at most one record per position
    uint32_t *tile_info = g->m3_tile_info.as<uint32_t>();
    uint32_t *H1T = g->status.as<uint32_t>(), *H12 = g->scratch1.as<uint32_t>();
    unsigned int *tot32 = g->hist.as<unsigned int>();
    uint32_t *base32 = g->m3_gbase.as<uint32_t>() + (size_t)gi * (cells + 1);   // kept for pass 3
    PDG_CUDA(cudaMemsetAsync(tot32, 0, cells * 4, s));
    const uint2 *packed = g->packed.as<uint2>();
    const uint64_t pw = g->packed.cap / 8;
    m3_tile_info_kernel<<<(pl.n_tiles + 255) / 256, 256, 0, s>>>(ct, pl, kv, tile_info);
    PDG_LAUNCHED(g);
    // CUDA events around every launch of the sort (bench.py's roofline numbers): event i precedes
    // stage i of {hist, scans, pass 1, pass 2, pass 3}
    const bool timed = gi == 0 && g->m3_max_groups == 1;   // a single group's launches run back to back on s
    if (timed)
        while (g->k_events.size() < 6) {
            cudaEvent_t ev;
            PDG_CUDA(cudaEventCreate(&ev));
            g->k_events.push_back(ev);
        }
    auto mark = [&](int i) { if (timed) PDG_CUDA(cudaEventRecord(g->k_events[i], s)); };
    mark(0);
    m3_hist_kernel<<<pl.n_chunks, M3H_THREADS, cells * 4, s>>>(packed, pw, k, kv, ct, pl, tile_info, H1T, H12, tot32);
    PDG_LAUNCHED(g);
    mark(1);
    uint32_t *goff = g->m3_gbase.as<uint32_t>() + (size_t)g->m3_max_groups * (cells + 1);
    m3_base_kernel<<<1, 1024, 0, s>>>(tot32, cells, base32, goff + gi, (uint32_t)rec_off);
    PDG_LAUNCHED(g);
    m3_scan_chunks_kernel<<<(cells + 255) / 256, 256, 0, s>>>(H12, pl.n_chunks, cells, base32);
    PDG_LAUNCHED(g);
    m3_scan_tiles_kernel<<<pl.n_chunks, 256, 0, s>>>(H1T, pl, H12, base32);
    PDG_LAUNCHED(g);
    mark(2);
    m3_pass1_kernel<7><<<(pl.n_tiles + M3_P1_TILES - 1) / M3_P1_TILES, M3_THREADS, sizeof(M3Smem2), s>>>(packed, pw, k, kv, ct, pl, tile_info, H1T,
                                                                     g->recA.as<uint64_t>() + rec_off);
    PDG_LAUNCHED(g);
    mark(3);
    m3_pass2_kernel<M3_B2><<<pl.n_chunks * B3, M3_THREADS, sizeof(M3Smem2), s>>>(g->recA.as<uint2>() + rec_off, pl, H1T, H12, base32,
                                                                          g->recB.as<uint64_t>() + rec_off,
                                                                          g->m3_use_d1 ? g->m3_d1.as<uint8_t>() + rec_off : nullptr);
    PDG_LAUNCHED(g);
    mark(4);
    g->m3_timed = timed;
    g->m3_groups.push_back(pdg_graph::M3Group{pos0, pos1, rec_off});
}

// Pass 3 + presence + ranks, all asynchronous on the graph's stream: a pipelined ingest launches
// this speculatively when its last group is staged, so that it runs while the caller declares
// the contigs.  msd3_finish_complete reads the node count back and writes the rows.
void msd3_finish_launch(pdg_graph *g, bool rows) {
    cudaStream_t s = g->stream;
    const int kv = g->k, k = std::min(g->k, kMsd3MaxBits / 2);   // sort key: the first k residues of a kv-mer
    (void)kv;
    const M3Plan pl = m3_plan(k, 0, 32);   // digit widths only
    const uint32_t cells = 1u << (pl.b3 + pl.b2);
    const uint32_t G = (uint32_t)g->m3_groups.size();
    const uint64_t n_codes = 1ULL << (2 * k);
    uint32_t *gbase = g->m3_gbase.as<uint32_t>();
    uint32_t *goff = gbase + (size_t)g->m3_max_groups * (cells + 1);   // written by every group's m3_base_kernel
    unsigned int *tot32 = g->hist.as<unsigned int>();
    uint32_t *base32 = g->hist.as<uint32_t>() + cells;
    uint32_t *rowtab = g->scratch0.as<uint32_t>();
    if (G == 1) {
        base32 = gbase;   // the group's bases are the global ones
    } else {
        m3_cell_totals_kernel<<<(cells + 255) / 256, 256, 0, s>>>(gbase, G, cells, tot32);
        PDG_LAUNCHED(g);
        m3_base_kernel<<<1, 1024, 0, s>>>(tot32, cells, base32, nullptr, 0u);
        PDG_LAUNCHED(g);
    }
    m3_pass3_kernel<M3_B1><<<cells, M3_THREADS, sizeof(M3Smem2), s>>>(g->recB.as<uint2>(), pl, gbase, goff, G, base32, rowtab,
                                                               g->g_pos.as<uint32_t>(),
                                                               g->m3_use_d1 ? g->m3_d1.as<uint8_t>() : nullptr);
    PDG_LAUNCHED(g);
    if (G == 1 && g->m3_timed)   // the group's other stages were timed
        PDG_CUDA(cudaEventRecord(g->k_events[5], s));
    // sentinel: rowtab[n_codes] = number of records (= base of the cell past the last)
    PDG_CUDA(cudaMemcpyAsync(rowtab + n_codes, base32 + cells, 4, cudaMemcpyDeviceToDevice, s));
    g->m3_launched = true;
    if (!rows) return;   // k > 11 (msdlong.cu): the buckets of the 11-residue prefix are finished there
    // rows: presence -> rank
    uint32_t *flags = g->recA.as<uint32_t>(), *rank = g->scratch1.as<uint32_t>();
    m3_present_kernel<<<grid_for(n_codes, 256), 256, 0, s>>>(rowtab, n_codes, flags);
    PDG_LAUNCHED(g);
    exclusive_scan_u32_u32(flags, rank, n_codes, g);
    g->m3_launched = true;
}

void msd3_finish_complete(pdg_graph *g, uint64_t n) {
    cudaStream_t s = g->stream;
    const int kv = g->k, k = std::min(g->k, kMsd3MaxBits / 2);   // sort key: the first k residues of a kv-mer
    (void)kv;
    const uint32_t G = (uint32_t)g->m3_groups.size();
    const uint64_t n_codes = 1ULL << (2 * k);
    uint32_t *rowtab = g->scratch0.as<uint32_t>();
    uint32_t *flags = g->recA.as<uint32_t>(), *rank = g->scratch1.as<uint32_t>();
    uint32_t nn = 0;
    PDG_CUDA(cudaMemcpyAsync(&nn, rank + n_codes, 4, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    if (G == 1 && g->m3_timed) {
        float ms[5] = {0, 0, 0, 0, 0};
        for (int i = 0; i < 5; ++i) cudaEventElapsedTime(&ms[i], g->k_events[i], g->k_events[i + 1]);
        g->stats.ms_hist = ms[0];
        g->stats.ms_scans = ms[1];
        g->stats.ms_pass[0] = ms[2];
        g->stats.ms_pass[1] = ms[3];
        g->stats.ms_pass[2] = ms[4];
        g->stats.ms_sort_kernel_avg = (ms[2] + ms[3] + ms[4]) / 3.0f;
        g->stats.sort_kernel_launches = 3;
    }
    g->stats.sort_passes = 3;
    g->stats.sort_records = n;
    g->stats.sort_groups = G;
    g->g_nodes = nn;
    g->g_codes.reserve(std::max<size_t>((size_t)nn * 8, 16), s);
    g->g_rowptr.reserve(((size_t)nn + 1) * 4, s);
    m3_compact_rows_kernel<<<grid_for(n_codes, 256), 256, 0, s>>>(rowtab, flags, rank, n_codes,
                                                                 g->g_codes.as<uint64_t>(), g->g_rowptr.as<uint32_t>());
    PDG_LAUNCHED(g);
    g->m3_groups.clear();
    g->m3_launched = false;
}

void msd3_finish(pdg_graph *g, uint64_t n) {
    if (!g->m3_launched) msd3_finish_launch(g, true);
    msd3_finish_complete(g, n);
}

void msd3_build(pdg_graph *g, const ContigTable &ct, const HostContigs &hc) {
    msd3_reserve(g, hc.pos0, hc.pos1, 1);
    msd3_stage_group(g, ct, hc.pos0, hc.pos1);
    msd3_finish(g, hc.n_kmers);
}

}  // namespace pdg
