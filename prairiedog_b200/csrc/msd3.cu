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

namespace pdg {

struct M3Plan {
    int b1, b2, b3;              // digit widths, low to high
    uint32_t n_tiles, ch, n_chunks;
    uint64_t pos0, pos1;
};

struct M3Contigs {
    uint64_t ci;
    bool simple;   // every window of the tile is a k-mer of one contig
};

__device__ __forceinline__ void m3_tile_contigs(const ContigTable &ct, uint64_t p0, uint64_t p1, int k, M3Contigs *tc) {
    const uint64_t ci = contig_of(ct, p0);
    tc->ci = ci;
    tc->simple = p0 >= ct.begin[ci] && p1 + k <= ct.end[ci] + 1;
}
__device__ __forceinline__ bool m3_valid(const ContigTable &ct, const M3Contigs &tc, uint64_t p, int k) {
    if (tc.simple) return true;
    uint64_t ci = tc.ci;
    while (ci + 1 < ct.n && ct.begin[ci + 1] <= p) ++ci;
    return p >= ct.begin[ci] && p + k <= ct.end[ci];
}

// ---- histograms from the packed genome --------------------------------------------------------
// H1T[d3][tile]; H12[chunk][d3 << b2 | d2]; tot32[cell] += chunk counts.
// One CTA per chunk; every WARP takes whole tiles on its own (private d3 counters, its own copy of
// the tile's packed words), so there is no block barrier inside the tile loop.
static constexpr int M3H_THREADS = 1024;
static constexpr int M3H_WARPS = M3H_THREADS / 32;
__global__ void __launch_bounds__(M3H_THREADS) m3_hist_kernel(const unsigned long long *__restrict__ packed,
                                                              uint64_t packed_words, int k, ContigTable ct, M3Plan pl,
                                                              uint32_t *__restrict__ H1T, uint32_t *__restrict__ H12,
                                                              unsigned int *__restrict__ tot32) {
    extern __shared__ uint32_t h12[];   // [cells]
    __shared__ uint32_t h1[M3H_WARPS][128];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t cells = 1u << (pl.b3 + pl.b2), B3 = 1u << pl.b3;
    const int sh2 = pl.b1, sh3 = pl.b1 + pl.b2;
    for (uint32_t i = tid; i < cells; i += M3H_THREADS) h12[i] = 0;
    __syncthreads();
    const uint32_t t0 = blockIdx.x * pl.ch, t1 = min(t0 + pl.ch, pl.n_tiles);
    for (uint32_t tile = t0 + w; tile < t1; tile += M3H_WARPS) {
        const uint64_t p0 = pl.pos0 + (uint64_t)tile * M3_TILE, p1 = min(p0 + (uint64_t)M3_TILE, pl.pos1);
        for (int i = lane; i < 128; i += 32) h1[w][i] = 0;
        M3Contigs tc;
        m3_tile_contigs(ct, p0, p1, k, &tc);
        __syncwarp();
        // lane l walks positions p0 + 32*i + l: the window starts in word (p >> 5) and may spill
        // into the next one, so keep a sliding pair of words
        const uint64_t w0 = p0 >> 5;
        const uint32_t lead = (uint32_t)(p0 & 31u);
        const uint32_t n_steps = (uint32_t)((p1 - p0 + 31) >> 5);
        for (uint32_t i = 0; i < n_steps; ++i) {
            const uint64_t q = (uint64_t)lead + 32u * i + lane;       // slot relative to word w0
            const uint64_t p = p0 + 32u * i + lane;
            const uint64_t wi = w0 + (q >> 5);
            const unsigned long long hi = wi < packed_words ? packed[wi] : 0ULL;
            const unsigned long long lo = wi + 1 < packed_words ? packed[wi + 1] : 0ULL;
            const uint32_t sft = 2u * (uint32_t)(q & 31u);
            const unsigned long long x = sft ? ((hi << sft) | (lo >> (64 - sft))) : hi;
            const uint32_t code = (uint32_t)(x >> (64 - 2 * k));
            if (p < p1 && m3_valid(ct, tc, p, k)) {
                atomicAdd(&h1[w][code >> sh3], 1u);
                atomicAdd(&h12[code >> sh2], 1u);
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
                                                       uint32_t *__restrict__ base32) {
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

// column d3 (one CTA): H1T[d3][tile] -> base3[d3] + exclusive prefix over tiles, in place
__global__ void __launch_bounds__(1024) m3_scan_tiles_kernel(uint32_t *__restrict__ H1T, uint32_t n_tiles,
                                                             const uint32_t *__restrict__ base32, int b2) {
    uint32_t *col = H1T + (size_t)blockIdx.x * n_tiles;
    uint64_t carry = base32[(size_t)blockIdx.x << b2];
    for (uint32_t b = 0; b < n_tiles; b += 1024) {
        const uint32_t i = b + threadIdx.x;
        const uint64_t v = i < n_tiles ? col[i] : 0;
        uint64_t tot;
        const uint64_t ex = block_excl_scan(v, &tot);
        if (i < n_tiles) col[i] = (uint32_t)(carry + ex);
        carry += tot;
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
__global__ void __launch_bounds__(M3_THREADS, 3) m3_pass1_kernel(const unsigned long long *__restrict__ packed,
                                                                 uint64_t packed_words, int k, ContigTable ct, M3Plan pl,
                                                                 const uint32_t *__restrict__ O1T,
                                                                 uint64_t *__restrict__ out) {
    extern __shared__ __align__(16) unsigned char m3_raw[];
    M3Smem &sm = *reinterpret_cast<M3Smem *>(m3_raw);
    __shared__ unsigned long long words[M3_WORDS];
    __shared__ M3Contigs tc;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t tile = blockIdx.x, B3 = 1u << pl.b3;
    const int sh3 = pl.b1 + pl.b2;
    const uint64_t p0 = pl.pos0 + (uint64_t)tile * M3_TILE, p1 = min(p0 + (uint64_t)M3_TILE, pl.pos1);
    for (int i = tid; i < M3_WORDS; i += M3_THREADS) {
        const uint64_t wi = (p0 >> 5) + i;
        words[i] = wi < packed_words ? packed[wi] : 0ULL;
    }
    if (tid < (int)B3) sm.cursor[tid] = O1T[(size_t)tid * pl.n_tiles + tile];
    if (tid == 0) m3_tile_contigs(ct, p0, p1, k, &tc);
    __syncthreads();
    const uint32_t lead = (uint32_t)(p0 & 31u);
    uint64_t r[M3_ITEMS];
    uint32_t dg[M3_ITEMS];
#pragma unroll
    for (int j = 0; j < M3_ITEMS; ++j) {
        const uint32_t li = w * (M3_ITEMS * 32) + j * 32 + lane;
        const uint64_t p = p0 + li;
        const uint32_t code = (uint32_t)kmer_at(words, lead + li, k);
        r[j] = ((uint64_t)code << 32) | (uint32_t)p;
        dg[j] = (p < p1 && m3_valid(ct, tc, p, k)) ? (code >> sh3) : M3_DROP;
    }
    m3_partition_tile<false>(sm, r, dg, sh3, (int)B3, out, nullptr);
}

// ---- pass 2: range (chunk, d3) -> grouped by (d3, d2) --------------------------------------------
__global__ void __launch_bounds__(M3_THREADS, 3) m3_pass2_kernel(const uint64_t *__restrict__ in, M3Plan pl,
                                                                 const uint32_t *__restrict__ O1T,
                                                                 const uint32_t *__restrict__ O2,
                                                                 const uint32_t *__restrict__ base32,
                                                                 uint64_t *__restrict__ out) {
    extern __shared__ __align__(16) unsigned char m3_raw[];
    M3Smem &sm = *reinterpret_cast<M3Smem *>(m3_raw);
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t B2 = 1u << pl.b2, B3 = 1u << pl.b3;
    const uint32_t c = blockIdx.x / B3, d3 = blockIdx.x % B3;
    const uint32_t begin = O1T[(size_t)d3 * pl.n_tiles + (size_t)c * pl.ch];
    const uint32_t end = (c + 1 < pl.n_chunks) ? O1T[(size_t)d3 * pl.n_tiles + (size_t)(c + 1) * pl.ch]
                                                : base32[(size_t)(d3 + 1) << pl.b2];
    if (begin >= end) return;
    if (tid < (int)B2) sm.cursor[tid] = O2[(size_t)c * (B3 * B2) + ((size_t)d3 << pl.b2) + tid];
    __syncthreads();
    for (uint32_t t0 = begin; t0 < end; t0 += M3_TILE) {
        const uint32_t tile_n = min((uint32_t)M3_TILE, end - t0);
        uint64_t r[M3_ITEMS];
        uint32_t dg[M3_ITEMS];
#pragma unroll
        for (int j = 0; j < M3_ITEMS; ++j) {
            const uint32_t li = w * (M3_ITEMS * 32) + j * 32 + lane;
            const bool ok = li < tile_n;
            r[j] = ok ? in[t0 + li] : 0ULL;
            dg[j] = ok ? (((uint32_t)(r[j] >> 32) >> pl.b1) & (B2 - 1)) : M3_DROP;
        }
        m3_partition_tile<false>(sm, r, dg, pl.b1, (int)B2, out, nullptr);
    }
}

// ---- pass 3: bucket (d3, d2): count d1, row starts, position column ------------------------------
__global__ void __launch_bounds__(M3_THREADS, 3) m3_pass3_kernel(const uint64_t *__restrict__ in, M3Plan pl,
                                                                 const uint32_t *__restrict__ base32,
                                                                 uint32_t *__restrict__ rowtab,
                                                                 uint32_t *__restrict__ out_pos) {
    extern __shared__ __align__(16) unsigned char m3_raw[];
    M3Smem &sm = *reinterpret_cast<M3Smem *>(m3_raw);
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t B1 = 1u << pl.b1, cell = blockIdx.x;
    const uint32_t begin = base32[cell], end = base32[cell + 1];
    // count d1 over the bucket (sm.bstart doubles as the histogram)
    if (tid < M3_MAX_BINS) sm.bstart[tid] = 0;
    __syncthreads();
    for (uint32_t i = begin + tid; i < end; i += M3_THREADS)
        atomicAdd(&sm.bstart[(uint32_t)(in[i] >> 32) & (B1 - 1)], 1u);
    __syncthreads();
    if (w == 0) {
        constexpr int PER = M3_MAX_BINS / 32;
        uint32_t v[PER], s = 0;
#pragma unroll
        for (int i = 0; i < PER; ++i) { v[i] = sm.bstart[lane * PER + i]; s += v[i]; }
        uint32_t inc = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        uint32_t ex = begin + inc - s;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            const uint32_t b = lane * PER + i;
            sm.cursor[b] = ex;
            if (b < B1) rowtab[((size_t)cell << pl.b1) + b] = ex;   // first occurrence of code (cell, b)
            ex += v[i];
        }
    }
    __syncthreads();
    for (uint32_t t0 = begin; t0 < end; t0 += M3_TILE) {
        const uint32_t tile_n = min((uint32_t)M3_TILE, end - t0);
        uint64_t r[M3_ITEMS];
        uint32_t dg[M3_ITEMS];
#pragma unroll
        for (int j = 0; j < M3_ITEMS; ++j) {
            const uint32_t li = w * (M3_ITEMS * 32) + j * 32 + lane;
            const bool ok = li < tile_n;
            r[j] = ok ? in[t0 + li] : 0ULL;
            dg[j] = ok ? ((uint32_t)(r[j] >> 32) & (B1 - 1)) : M3_DROP;
        }
        m3_partition_tile<true>(sm, r, dg, 0, (int)B1, nullptr, out_pos);
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

void msd3_build(pdg_graph *g, const ContigTable &ct, const HostContigs &hc) {
    cudaStream_t s = g->stream;
    const int k = g->k;
    const uint64_t n = hc.n_kmers, span = hc.pos1 - hc.pos0;
    M3Plan pl{};
    pl.b1 = std::min(8, 2 * k);
    pl.b2 = std::min(7, 2 * k - pl.b1);
    pl.b3 = 2 * k - pl.b1 - pl.b2;
    pl.pos0 = hc.pos0;
    pl.pos1 = hc.pos1;
    pl.n_tiles = (uint32_t)((span + M3_TILE - 1) / M3_TILE);
    const uint32_t B3 = 1u << pl.b3, cells = 1u << (pl.b3 + pl.b2);
    pl.ch = std::max<uint32_t>(1, (3500u * B3) / M3_TILE);   // ranges of pass 2 average ~3500 records
    pl.n_chunks = (pl.n_tiles + pl.ch - 1) / pl.ch;
    const uint64_t n_codes = 1ULL << (2 * k);

    g->status.reserve((size_t)B3 * pl.n_tiles * 4, s);                 // H1T -> O1T
    g->scratch1.reserve(std::max<size_t>((size_t)pl.n_chunks * cells * 4, (n_codes + 1) * 4), s);   // H12 -> O2, later ranks
    g->hist.reserve((size_t)(2 * cells + 2) * 4, s);                    // tot32 | base32
    g->scratch0.reserve((n_codes + 1) * 4, s);                          // row starts of every code
    g->recA.reserve(std::max<size_t>(n * 8, n_codes * 4), s);
    g->recB.reserve(n * 8, s);
    g->g_pos.reserve(n * 4, s);
    uint32_t *H1T = g->status.as<uint32_t>(), *H12 = g->scratch1.as<uint32_t>();
    unsigned int *tot32 = g->hist.as<unsigned int>();
    uint32_t *base32 = g->hist.as<uint32_t>() + cells;
    uint32_t *rowtab = g->scratch0.as<uint32_t>();
    PDG_CUDA(cudaMemsetAsync(tot32, 0, cells * 4, s));

    static bool attr_set = false;
    if (!attr_set) {
        PDG_CUDA(cudaFuncSetAttribute(m3_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 1 << 16));
        PDG_CUDA(cudaFuncSetAttribute(m3_pass1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(M3Smem)));
        PDG_CUDA(cudaFuncSetAttribute(m3_pass2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(M3Smem)));
        PDG_CUDA(cudaFuncSetAttribute(m3_pass3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(M3Smem)));
        attr_set = true;
    }
    const unsigned long long *packed = g->packed.as<unsigned long long>();
    const uint64_t pw = g->packed.cap / 8;
    m3_hist_kernel<<<pl.n_chunks, M3H_THREADS, cells * 4, s>>>(packed, pw, k, ct, pl, H1T, H12, tot32);
    PDG_LAUNCHED(g);
    m3_base_kernel<<<1, 1024, 0, s>>>(tot32, cells, base32);
    PDG_LAUNCHED(g);
    m3_scan_tiles_kernel<<<B3, 1024, 0, s>>>(H1T, pl.n_tiles, base32, pl.b2);
    PDG_LAUNCHED(g);
    m3_scan_chunks_kernel<<<(cells + 255) / 256, 256, 0, s>>>(H12, pl.n_chunks, cells, base32);
    PDG_LAUNCHED(g);

    const bool timed = g->collect_sort_stats;
    if (timed && !g->ev0) {
        PDG_CUDA(cudaEventCreate(&g->ev0));
        PDG_CUDA(cudaEventCreate(&g->ev1));
    }
    if (timed) PDG_CUDA(cudaEventRecord(g->ev0, s));
    m3_pass1_kernel<<<pl.n_tiles, M3_THREADS, sizeof(M3Smem), s>>>(packed, pw, k, ct, pl, H1T, g->recA.as<uint64_t>());
    PDG_LAUNCHED(g);
    m3_pass2_kernel<<<pl.n_chunks * B3, M3_THREADS, sizeof(M3Smem), s>>>(g->recA.as<uint64_t>(), pl, H1T, H12, base32,
                                                                         g->recB.as<uint64_t>());
    PDG_LAUNCHED(g);
    m3_pass3_kernel<<<cells, M3_THREADS, sizeof(M3Smem), s>>>(g->recB.as<uint64_t>(), pl, base32, rowtab,
                                                              g->g_pos.as<uint32_t>());
    PDG_LAUNCHED(g);
    if (timed) {
        PDG_CUDA(cudaEventRecord(g->ev1, s));
        PDG_CUDA(cudaEventSynchronize(g->ev1));
        float ms = 0;
        cudaEventElapsedTime(&ms, g->ev0, g->ev1);
        g->stats.ms_sort_kernel_avg = ms / 3.0f;
        g->stats.sort_kernel_launches = 3;
    }
    g->stats.sort_passes = 3;
    g->stats.sort_records = n;
    // sentinel: rowtab[n_codes] = n
    const uint32_t n32 = (uint32_t)n;
    PDG_CUDA(cudaMemcpyAsync(rowtab + n_codes, &n32, 4, cudaMemcpyHostToDevice, s));
    // rows: presence -> rank -> (codes, rowptr)
    uint32_t *flags = g->recA.as<uint32_t>(), *rank = g->scratch1.as<uint32_t>();
    m3_present_kernel<<<grid_for(n_codes, 256), 256, 0, s>>>(rowtab, n_codes, flags);
    PDG_LAUNCHED(g);
    exclusive_scan_u32_u32(flags, rank, n_codes, g);
    uint32_t nn = 0;
    PDG_CUDA(cudaMemcpyAsync(&nn, rank + n_codes, 4, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    g->g_nodes = nn;
    g->g_codes.reserve(std::max<size_t>((size_t)nn * 8, 16), s);
    g->g_rowptr.reserve(((size_t)nn + 1) * 4, s);
    m3_compact_rows_kernel<<<grid_for(n_codes, 256), 256, 0, s>>>(rowtab, flags, rank, n_codes,
                                                                 g->g_codes.as<uint64_t>(), g->g_rowptr.as<uint32_t>());
    PDG_LAUNCHED(g);
}

}  // namespace pdg
