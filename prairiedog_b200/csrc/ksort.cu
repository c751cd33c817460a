// ksort.cu -- the occurrence-index sort for k <= 16, fused with k-mer emission.
//
// (code, position) records are never materialised in emission order: the first partition pass
// reads the 2-bit packed genome directly (0.25 B per base), forms each window's code in
// registers and scatters (code << 32 | position) by the least significant digit; middle passes
// move 8-byte records; the last pass writes only the 4-byte position column and, for k <= 13,
// finds every row start with an atomicMin into a direct table indexed by code.  All digit
// histograms come from one pass over the packed genome; bin bases are scanned on the device, so
// the whole sort runs without a host round trip.
//
// Each pass is one onesweep-style kernel (a CTA takes a tile ticket, ranks its records with
// warp match + one shared-memory atomic per distinct digit, learns its global bin offsets by
// decoupled look-back, stages the tile in sorted order in shared memory and writes runs with
// coalesced stores).  HBM traffic per record: first pass 0.25 + 8 B, middle 8 + 8 B, last 8 + 4 B.
//
// Reference counterpart: the per-key B-tree inserts of lemongraph/lib/lemongraph.c:591-630
// (node_idx / srcnode_idx / tgtnode_idx) that keep LemonGraph's indexes ordered.
#include "gindex.cuh"
#include "ksort.cuh"
#include <algorithm>

namespace pdg {

static constexpr unsigned long long ST_LOCAL = 1ULL << 62;
static constexpr unsigned long long ST_GLOBAL = 1ULL << 63;
static constexpr unsigned long long ST_VALUE = (1ULL << 62) - 1;

// validity of the window starting at position p: inside a contig with p + k <= end
struct TileContigs {
    uint64_t cb, ce;   // contig that contains the tile start (or the first one after it)
    uint64_t ci;
    bool simple;       // every position of the tile lies in [cb, ce - k]
};

__device__ __forceinline__ bool window_valid(const ContigTable &ct, const TileContigs &tc, uint64_t p, int k) {
    if (tc.simple) return true;
    uint64_t ci = tc.ci;
    while (ci + 1 < ct.n && ct.begin[ci + 1] <= p) ++ci;
    return p >= ct.begin[ci] && p + k <= ct.end[ci];
}

__device__ __forceinline__ void tile_contigs(const ContigTable &ct, uint64_t p0, uint64_t p1, int k, TileContigs *tc) {
    const uint64_t ci = contig_of(ct, p0);
    tc->ci = ci;
    tc->cb = ct.begin[ci];
    tc->ce = ct.end[ci];
    tc->simple = p0 >= tc->cb && p1 + k <= tc->ce + 1;   // last position p1-1: p1-1+k <= ce
}

// ---- histograms of every pass digit, straight from the packed genome -------------------------
__global__ void __launch_bounds__(KS_THREADS) ksort_hist_kernel(const unsigned long long *__restrict__ packed,
                                                                uint64_t packed_words, uint64_t pos0,
                                                                uint64_t pos1, int k, ContigTable ct, KsPasses pl,
                                                                unsigned long long *__restrict__ hist) {
    __shared__ unsigned long long words[KS_TILE / 32 + 2];
    __shared__ uint32_t sh[KS_MAX_PASSES * 256];
    __shared__ TileContigs tc;
    for (int i = threadIdx.x; i < pl.n * 256; i += KS_THREADS) sh[i] = 0;
    const uint64_t n_tiles = (pos1 - pos0 + KS_TILE - 1) / KS_TILE;
    for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint64_t p0 = pos0 + tile * KS_TILE, p1 = min(p0 + (uint64_t)KS_TILE, pos1);
        const uint64_t w0 = p0 >> 5;
        __syncthreads();
        for (int i = threadIdx.x; i < KS_TILE / 32 + 2; i += KS_THREADS)
            words[i] = (w0 + i < packed_words) ? packed[w0 + i] : 0ULL;
        if (threadIdx.x == 0) tile_contigs(ct, p0, p1, k, &tc);
        __syncthreads();
        const uint32_t lead = (uint32_t)(p0 & 31u);
#pragma unroll
        for (int j = 0; j < KS_ITEMS; ++j) {
            const uint32_t l = j * KS_THREADS + threadIdx.x;
            const uint64_t p = p0 + l;
            if (p < p1 && window_valid(ct, tc, p, k)) {
                const uint32_t code = (uint32_t)kmer_at(words, lead + l, k);
                for (int q = 0; q < pl.n; ++q)
                    atomicAdd(&sh[q * 256 + ((code >> pl.shift[q]) & pl.mask[q])], 1u);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < pl.n * 256; i += KS_THREADS)
        if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}

// hist[p][256] -> base[p][256] (exclusive scan per pass); one CTA of 256 threads per pass
__global__ void __launch_bounds__(256) ksort_bases_kernel(const unsigned long long *__restrict__ hist,
                                                          uint64_t *__restrict__ base) {
    const uint64_t v = hist[blockIdx.x * 256 + threadIdx.x];
    base[blockIdx.x * 256 + threadIdx.x] = block_excl_scan(v, nullptr);
}

// ---- one partition pass -----------------------------------------------------------------------
// IN_GENOME: items are the windows of positions [pos0, pos1) (invalid windows are dropped);
//            otherwise items are the n records of `in`.
// OUT_FINAL: writes the position column and row starts instead of records.
template <bool IN_GENOME, bool OUT_FINAL>
__global__ void __launch_bounds__(KS_THREADS, 2)
ksort_pass_kernel(const unsigned long long *__restrict__ packed, uint64_t packed_words, uint64_t pos0, uint64_t pos1,
                  int k, ContigTable ct, const uint64_t *__restrict__ in, uint64_t n, int shift, uint32_t mask,
                  const uint64_t *__restrict__ bin_base, volatile unsigned long long *status, unsigned int *ticket,
                  uint64_t *__restrict__ out_rec, uint32_t *__restrict__ out_pos, uint32_t *__restrict__ row_start) {
    extern __shared__ __align__(16) unsigned char ks_smem[];
    uint64_t *stage = reinterpret_cast<uint64_t *>(ks_smem);                              // [KS_TILE]
    uint32_t *cnt = reinterpret_cast<uint32_t *>(ks_smem + sizeof(uint64_t) * KS_TILE);   // [KS_WARPS][256]
    __shared__ uint32_t bin_start[256];
    __shared__ uint64_t gbase[256];
    __shared__ unsigned long long words[IN_GENOME ? KS_TILE / 32 + 2 : 1];
    __shared__ uint32_t tile_s;
    __shared__ TileContigs tc;

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) tile_s = atomicAdd(ticket, 1u);
    for (int i = tid; i < KS_WARPS * 256; i += KS_THREADS) cnt[i] = 0;
    __syncthreads();
    const uint64_t tile = tile_s;
    const uint64_t item0 = tile * KS_TILE;
    const uint64_t n_items = IN_GENOME ? (pos1 - pos0) : n;
    const uint32_t tile_n = (uint32_t)min((uint64_t)KS_TILE, n_items - item0);

    // 1. load (warp-striped: warp w owns items [w*256, (w+1)*256) of the tile)
    uint64_t r[KS_ITEMS];
    uint32_t dig[KS_ITEMS];   // digit, or 0xFFFFFFFF for a dropped item
    const uint32_t wbase = w * (KS_ITEMS * 32);
    if (IN_GENOME) {
        const uint64_t p0 = pos0 + item0;
        const uint64_t w0 = p0 >> 5;
        for (int i = tid; i < KS_TILE / 32 + 2; i += KS_THREADS)
            words[i] = (w0 + i < packed_words) ? packed[w0 + i] : 0ULL;
        if (tid == 0) tile_contigs(ct, p0, p0 + tile_n, k, &tc);
        __syncthreads();
        const uint32_t lead = (uint32_t)(p0 & 31u);
#pragma unroll
        for (int j = 0; j < KS_ITEMS; ++j) {
            const uint32_t li = wbase + j * 32 + lane;
            const uint64_t p = p0 + li;
            const uint32_t code = (uint32_t)kmer_at(words, lead + li, k);
            r[j] = ((uint64_t)code << 32) | (uint32_t)p;
            const bool ok = li < tile_n && window_valid(ct, tc, p, k);
            dig[j] = ok ? ((code >> shift) & mask) : 0xFFFFFFFFu;
        }
    } else {
#pragma unroll
        for (int j = 0; j < KS_ITEMS; ++j) {
            const uint32_t li = wbase + j * 32 + lane;
            const bool ok = li < tile_n;
            r[j] = ok ? in[item0 + li] : 0ULL;
            dig[j] = ok ? ((uint32_t)(r[j] >> (32 + shift)) & mask) : 0xFFFFFFFFu;
        }
    }

    // 2. stable rank inside the warp: all matches first, then one ordered atomic per distinct
    //    digit of each round, then the broadcasts (three independent instruction streams)
    uint32_t *wcnt = cnt + w * 256;
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint32_t peers[KS_ITEMS], rank[KS_ITEMS];
#pragma unroll
    for (int j = 0; j < KS_ITEMS; ++j) peers[j] = __match_any_sync(0xffffffffu, dig[j]);
#pragma unroll
    for (int j = 0; j < KS_ITEMS; ++j) {
        const int leader = __ffs(peers[j]) - 1;
        uint32_t old = 0;
        if (lane == leader && dig[j] != 0xFFFFFFFFu) old = atomicAdd(&wcnt[dig[j]], (uint32_t)__popc(peers[j]));
        rank[j] = old;
    }
#pragma unroll
    for (int j = 0; j < KS_ITEMS; ++j) {
        const int leader = __ffs(peers[j]) - 1;
        rank[j] = __shfl_sync(0xffffffffu, rank[j], leader) + __popc(peers[j] & lt_mask);
    }
    __syncthreads();

    // 3. thread b < 256 owns bin b: exclusive offsets of the warps inside the bin, tile count,
    //    status for the tiles behind us, start of the bin inside the tile
    uint32_t tile_count = 0;
    if (tid < 256) {
#pragma unroll
        for (int ww = 0; ww < KS_WARPS; ++ww) {
            const uint32_t c = cnt[ww * 256 + tid];
            cnt[ww * 256 + tid] = tile_count;
            tile_count += c;
        }
        status[tile * 256 + tid] = (tile == 0 ? ST_GLOBAL : ST_LOCAL) | (unsigned long long)tile_count;
    }
    uint64_t total64;
    const uint32_t bstart = (uint32_t)block_excl_scan((uint64_t)tile_count, &total64);
    if (tid < 256) bin_start[tid] = bstart;
    __syncthreads();

    // 4. stage the tile in sorted order
#pragma unroll
    for (int j = 0; j < KS_ITEMS; ++j)
        if (dig[j] != 0xFFFFFFFFu) stage[bin_start[dig[j]] + wcnt[dig[j]] + rank[j]] = r[j];

    // 5. decoupled look-back: records of bin b in the tiles before this one
    if (tid < 256) {
        uint64_t excl = 0;
        if (tile != 0) {
            uint64_t t = tile - 1;
            while (true) {
                const unsigned long long v = status[t * 256 + tid];
                if ((v & (ST_LOCAL | ST_GLOBAL)) == 0) continue;   // predecessor not there yet
                excl += v & ST_VALUE;
                if (v & ST_GLOBAL) break;
                --t;   // tile 0 always posts GLOBAL, so t never underflows
            }
            status[tile * 256 + tid] = ST_GLOBAL | (unsigned long long)(excl + tile_count);
        }
        gbase[tid] = bin_base[tid] + excl;
    }
    __syncthreads();

    // 6. runs of equal digits go to consecutive global addresses
    const uint32_t total = (uint32_t)total64;
    for (uint32_t s = tid; s < total; s += KS_THREADS) {
        const uint64_t x = stage[s];
        const uint32_t d = (uint32_t)(x >> (32 + shift)) & mask;
        const uint64_t gi = gbase[d] + (s - bin_start[d]);
        if (OUT_FINAL) {
            out_pos[gi] = (uint32_t)x;
            if (row_start) {
                const uint32_t code = (uint32_t)(x >> 32);
                if (s == bin_start[d] || (uint32_t)(stage[s - 1] >> 32) != code) atomicMin(&row_start[code], (uint32_t)gi);
            }
        } else {
            out_rec[gi] = x;
        }
    }
}


// ---- host driver ------------------------------------------------------------------------------
__global__ void fill_u32_kernel(uint32_t *p, uint64_t n, uint32_t v) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// row_start[c] = first index of code c or 0xFFFFFFFF.  Emits presence flags for the scan.
__global__ void ks_present_kernel(const uint32_t *__restrict__ row_start, uint64_t n_codes, uint32_t *__restrict__ flags) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_codes) flags[i] = row_start[i] != 0xFFFFFFFFu ? 1u : 0u;
}
__global__ void ks_compact_rows_kernel(const uint32_t *__restrict__ row_start, const uint32_t *__restrict__ rank,
                                       uint64_t n_codes, uint64_t n_kmers, uint64_t *__restrict__ codes,
                                       uint32_t *__restrict__ rowptr) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_codes) {
        const uint32_t rs = row_start[i];
        if (rs != 0xFFFFFFFFu) {
            codes[rank[i]] = i;
            rowptr[rank[i]] = rs;
        }
    }
    if (i == 0) rowptr[rank[n_codes]] = (uint32_t)n_kmers;
}

static KsPasses plan_passes(int key_bits) {
    KsPasses pl{};
    pl.n = (key_bits + 7) / 8;
    int done = 0;
    for (int p = 0; p < pl.n; ++p) {
        const int bits = (key_bits - done + (pl.n - p) - 1) / (pl.n - p);   // spread evenly, wider first
        pl.shift[p] = done;
        pl.mask[p] = (1u << bits) - 1u;
        done += bits;
    }
    return pl;
}

template <bool IN_GENOME, bool OUT_FINAL>
static void launch_pass(pdg_graph *g, const ContigTable &ct, const HostContigs &hc, const uint64_t *in, uint64_t n,
                        int shift, uint32_t mask, const uint64_t *base, uint64_t n_tiles, unsigned int *ticket,
                        uint64_t *out_rec, uint32_t *out_pos, uint32_t *row_start) {
    const size_t smem = sizeof(uint64_t) * KS_TILE + sizeof(uint32_t) * KS_WARPS * 256;
    const uint32_t which = (uint32_t)pdg_graph::ATTR_KSORT << ((IN_GENOME ? 1 : 0) + (OUT_FINAL ? 2 : 0));   // one bit per instantiation
    if (!(g->attr_mask & which)) {
        PDG_CUDA(cudaFuncSetAttribute(ksort_pass_kernel<IN_GENOME, OUT_FINAL>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        g->attr_mask |= which;
    }
    ksort_pass_kernel<IN_GENOME, OUT_FINAL><<<(unsigned)n_tiles, KS_THREADS, smem, g->stream>>>(
        g->packed.as<unsigned long long>(), g->packed.cap / 8, hc.pos0, hc.pos1, g->k, ct, in, n, shift, mask, base,
        g->status.as<unsigned long long>(), ticket, out_rec, out_pos, row_start);
    PDG_LAUNCHED(g);
}

// Builds g_pos (+ g_codes / g_rowptr when the direct row table applies).  Returns false when the
// caller still has to derive the rows from the sorted records left in *sorted_out (13 < k <= 16).
bool ksort_build(pdg_graph *g, const ContigTable &ct, const HostContigs &hc, const uint64_t **sorted_out) {
    cudaStream_t s = g->stream;
    const int k = g->k;
    const uint64_t n = hc.n_kmers, span = hc.pos1 - hc.pos0;
    const bool direct = 2 * k <= kDirectMaxBits;
    const KsPasses pl = plan_passes(2 * k);
    const uint64_t tiles_g = (span + KS_TILE - 1) / KS_TILE, tiles_r = (n + KS_TILE - 1) / KS_TILE;
    // hist [np][256] u64 | base [np][256] u64 | tickets [np] u32
    const size_t hist_bytes = (size_t)pl.n * 256 * 8;
    g->hist.reserve(2 * hist_bytes + 256, s);
    unsigned long long *d_hist = g->hist.as<unsigned long long>();
    uint64_t *d_base = g->hist.as<uint64_t>() + (size_t)pl.n * 256;
    unsigned int *d_ticket = reinterpret_cast<unsigned int *>(g->hist.as<unsigned char>() + 2 * hist_bytes);
    PDG_CUDA(cudaMemsetAsync(g->hist.p, 0, 2 * hist_bytes + 256, s));
    const size_t status_bytes = std::max(tiles_g, tiles_r) * 256 * sizeof(unsigned long long);
    g->status.reserve(status_bytes, s);
    g->g_pos.reserve(n * 4, s);
    if (pl.n > 1) g->recA.reserve(n * 8, s);
    if (pl.n > 2 || (!direct && pl.n > 1)) g->recB.reserve(n * 8, s);
    if (!direct && pl.n == 1) g->recA.reserve(n * 8, s);
    uint32_t *row_start = nullptr;
    const uint64_t n_codes = 1ULL << (2 * k);
    if (direct) {
        g->scratch0.reserve((n_codes + 1) * 4, s);
        row_start = g->scratch0.as<uint32_t>();
        fill_u32_kernel<<<grid_for(n_codes, 256), 256, 0, s>>>(row_start, n_codes, 0xFFFFFFFFu);
        PDG_LAUNCHED(g);
    }
    const unsigned hist_grid = (unsigned)std::min<uint64_t>(tiles_g, 148 * 4);
    ksort_hist_kernel<<<hist_grid, KS_THREADS, 0, s>>>(g->packed.as<unsigned long long>(), g->packed.cap / 8, hc.pos0,
                                                       hc.pos1, k, ct, pl, d_hist);
    PDG_LAUNCHED(g);
    ksort_bases_kernel<<<pl.n, 256, 0, s>>>(d_hist, d_base);
    PDG_LAUNCHED(g);

    if (g->collect_sort_stats && !g->ev0) {
        PDG_CUDA(cudaEventCreate(&g->ev0));
        PDG_CUDA(cudaEventCreate(&g->ev1));
    }
    if (g->collect_sort_stats) PDG_CUDA(cudaEventRecord(g->ev0, s));
    uint64_t *bufs[2] = {g->recA.as<uint64_t>(), g->recB.as<uint64_t>()};
    const uint64_t *src = nullptr;
    for (int p = 0; p < pl.n; ++p) {
        const bool first = p == 0, last = p == pl.n - 1, fin = last && direct;
        const uint64_t n_tiles = first ? tiles_g : tiles_r;
        PDG_CUDA(cudaMemsetAsync(g->status.p, 0, n_tiles * 256 * sizeof(unsigned long long), s));
        uint64_t *dst = bufs[p & 1];
        const uint64_t *base = d_base + (size_t)p * 256;
        if (first && fin) launch_pass<true, true>(g, ct, hc, nullptr, n, pl.shift[p], pl.mask[p], base, n_tiles, d_ticket + p, nullptr, g->g_pos.as<uint32_t>(), row_start);
        else if (first) launch_pass<true, false>(g, ct, hc, nullptr, n, pl.shift[p], pl.mask[p], base, n_tiles, d_ticket + p, dst, nullptr, nullptr);
        else if (fin) launch_pass<false, true>(g, ct, hc, src, n, pl.shift[p], pl.mask[p], base, n_tiles, d_ticket + p, nullptr, g->g_pos.as<uint32_t>(), row_start);
        else launch_pass<false, false>(g, ct, hc, src, n, pl.shift[p], pl.mask[p], base, n_tiles, d_ticket + p, dst, nullptr, nullptr);
        src = dst;
    }
    if (g->collect_sort_stats) {
        PDG_CUDA(cudaEventRecord(g->ev1, s));
        PDG_CUDA(cudaEventSynchronize(g->ev1));
        float ms = 0;
        cudaEventElapsedTime(&ms, g->ev0, g->ev1);
        g->stats.ms_sort_kernel_avg = ms / (float)pl.n;   // includes the status memsets
        g->stats.sort_kernel_launches = pl.n;
    }
    g->stats.sort_passes = pl.n;
    g->stats.sort_records = n;
    if (!direct) {
        *sorted_out = src;
        return false;
    }
    // rows from the direct table: presence -> rank -> (codes, rowptr)
    g->scratch1.reserve((n_codes + 1) * 4, s);
    g->scratch2.reserve(16, s);
    DevBuf &flg = g->recB;   // free again: reuse as flags
    flg.reserve(n_codes * 4, s);
    ks_present_kernel<<<grid_for(n_codes, 256), 256, 0, s>>>(row_start, n_codes, flg.as<uint32_t>());
    PDG_LAUNCHED(g);
    exclusive_scan_u32_u32(flg.as<uint32_t>(), g->scratch1.as<uint32_t>(), n_codes, g);
    uint32_t nn = 0;
    PDG_CUDA(cudaMemcpyAsync(&nn, g->scratch1.as<uint32_t>() + n_codes, 4, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    g->g_nodes = nn;
    g->g_codes.reserve(std::max<size_t>((size_t)nn * 8, 16), s);
    g->g_rowptr.reserve(((size_t)nn + 1) * 4, s);
    ks_compact_rows_kernel<<<grid_for(n_codes, 256), 256, 0, s>>>(row_start, g->scratch1.as<uint32_t>(), n_codes, n,
                                                                 g->g_codes.as<uint64_t>(), g->g_rowptr.as<uint32_t>());
    PDG_LAUNCHED(g);
    return true;
}

}  // namespace pdg
