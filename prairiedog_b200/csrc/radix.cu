// radix.cu -- onesweep-style least-significant-digit radix sort (8-bit digits) for sm_100a.
//
// One histogram launch computes the digit histograms of every requested pass; each pass is then
// ONE kernel: a CTA takes a tile ticket, ranks its 4096 records with warp match/ballot, learns
// its global bin offsets by decoupled look-back over the tiles before it, stages the tile in
// shared memory in sorted order and writes runs of equal digits with coalesced stores.
// HBM traffic per pass: read n records + write n records (+ 2 KiB of status words per tile).
//
// Replaces the B-tree inserts of lemongraph/lib/lemongraph.c:591-630 (_node_index/_edge_index):
// the reference keeps its indexes ordered by inserting one key at a time into LMDB; here the
// same orderings (by node key, by (src,type,value), by (tgt,type,value)) come from bulk sorts.
#include "common.cuh"

namespace pdg {

static constexpr int RS_THREADS = 256;
static constexpr int RS_WARPS = RS_THREADS / 32;
static constexpr int RS_ITEMS = 16;
static constexpr int RS_TILE = RS_THREADS * RS_ITEMS;   // 4096 records
static constexpr int RS_MAX_PASSES = 16;

static constexpr unsigned long long ST_LOCAL = 1ULL << 62;
static constexpr unsigned long long ST_GLOBAL = 1ULL << 63;
static constexpr unsigned long long ST_VALUE = (1ULL << 62) - 1;

__device__ __forceinline__ uint32_t rec_field(const uint64_t &r, int field, int shift) {
    return (uint32_t)(r >> shift) & 0xFFu;
}
__device__ __forceinline__ uint32_t rec_field(const uint4 &r, int field, int shift) {
    uint32_t v = field == 0 ? r.x : field == 1 ? r.y : field == 2 ? r.z : r.w;
    return (v >> shift) & 0xFFu;
}

struct PassList {
    int n;
    int field[RS_MAX_PASSES];
    int shift[RS_MAX_PASSES];
};

// ---- histogram of every pass's digit, one read of the input --------------------------------
template <typename Rec>
__global__ void __launch_bounds__(RS_THREADS) radix_hist_kernel(const Rec *__restrict__ in, uint64_t n,
                                                                PassList pl,
                                                                unsigned long long *__restrict__ hist) {
    __shared__ uint32_t sh[RS_MAX_PASSES * 256];
    for (int i = threadIdx.x; i < pl.n * 256; i += RS_THREADS) sh[i] = 0;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
    for (uint64_t i = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x; i < n; i += stride) {
        const Rec r = in[i];
#pragma unroll 4
        for (int p = 0; p < pl.n; ++p) atomicAdd(&sh[p * 256 + rec_field(r, pl.field[p], pl.shift[p])], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < pl.n * 256; i += RS_THREADS)
        if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}

// ---- one onesweep pass ----------------------------------------------------------------------
template <typename Rec>
__global__ void __launch_bounds__(RS_THREADS, 2)
onesweep_kernel(const Rec *__restrict__ in, Rec *__restrict__ out, uint64_t n, int field, int shift,
                const uint64_t *__restrict__ bin_base,        // [256] start of each bin in `out`
                volatile unsigned long long *status,          // [n_tiles][256], zeroed
                unsigned int *ticket) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Rec *stage = reinterpret_cast<Rec *>(smem_raw);                                  // RS_TILE
    uint32_t *cnt = reinterpret_cast<uint32_t *>(smem_raw + sizeof(Rec) * RS_TILE);  // [warps][256]
    __shared__ uint32_t bin_start[256];
    __shared__ uint64_t gbase[256];
    __shared__ uint32_t tile_s;

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) tile_s = atomicAdd(ticket, 1u);
    for (int i = tid; i < RS_WARPS * 256; i += RS_THREADS) cnt[i] = 0;
    __syncthreads();
    const uint64_t tile = tile_s;
    const uint64_t tile_base = tile * RS_TILE;
    const uint32_t tile_n = (uint32_t)min((uint64_t)RS_TILE, n - tile_base);

    // 1. coalesced, warp-striped load: warp w owns records [w*512, (w+1)*512) of the tile
    Rec r[RS_ITEMS];
    uint32_t rank[RS_ITEMS];
    const uint32_t wbase = w * (RS_ITEMS * 32);
#pragma unroll
    for (int j = 0; j < RS_ITEMS; ++j) {
        const uint32_t li = wbase + j * 32 + lane;
        if (li < tile_n) r[j] = in[tile_base + li];
    }

    // 2. stable rank of every record among the records of its warp that share its digit
    uint32_t *wcnt = cnt + w * 256;
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int j = 0; j < RS_ITEMS; ++j) {
        const uint32_t li = wbase + j * 32 + lane;
        const bool valid = li < tile_n;
        const uint32_t d = valid ? rec_field(r[j], field, shift) : 256u;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (valid && lane == leader) {
            old = wcnt[d];
            wcnt[d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[j] = old + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();

    // 3. thread b owns bin b: exclusive offsets of the warps inside the bin, tile count
    uint32_t tile_count = 0;
#pragma unroll
    for (int ww = 0; ww < RS_WARPS; ++ww) {
        const uint32_t c = cnt[ww * 256 + tid];
        cnt[ww * 256 + tid] = tile_count;
        tile_count += c;
    }
    // 4. decoupled look-back: how many records of bin b precede this tile
    uint64_t excl = 0;
    if (tile == 0) {
        status[tid] = ST_GLOBAL | (unsigned long long)tile_count;
    } else {
        status[tile * 256 + tid] = ST_LOCAL | (unsigned long long)tile_count;
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
    const uint32_t bstart = (uint32_t)block_excl_scan((uint64_t)tile_count, nullptr);
    bin_start[tid] = bstart;
    __syncthreads();

    // 5. stage the tile in sorted order
#pragma unroll
    for (int j = 0; j < RS_ITEMS; ++j) {
        const uint32_t li = wbase + j * 32 + lane;
        if (li < tile_n) {
            const uint32_t d = rec_field(r[j], field, shift);
            stage[bin_start[d] + wcnt[d] + rank[j]] = r[j];
        }
    }
    __syncthreads();

    // 6. runs of equal digits go to consecutive global addresses
    for (uint32_t s = tid; s < tile_n; s += RS_THREADS) {
        const Rec x = stage[s];
        const uint32_t d = rec_field(x, field, shift);
        out[gbase[d] + (s - bin_start[d])] = x;
    }
}

template <typename Rec>
static Rec *radix_sort_impl(pdg_graph *g, Rec *a, Rec *b, uint64_t n, const std::vector<PassDesc> &passes,
                            uint64_t *passes_run) {
    cudaStream_t s = g->stream;
    if (passes_run) *passes_run = 0;
    if (n <= 1 || passes.empty()) return a;
    if (passes.size() > (size_t)RS_MAX_PASSES) {
        // split into groups (stable LSD: earlier groups are less significant)
        std::vector<PassDesc> first(passes.begin(), passes.begin() + RS_MAX_PASSES);
        std::vector<PassDesc> rest(passes.begin() + RS_MAX_PASSES, passes.end());
        uint64_t p1 = 0, p2 = 0;
        Rec *mid = radix_sort_impl<Rec>(g, a, b, n, first, &p1);
        Rec *other = mid == a ? b : a;
        Rec *res = radix_sort_impl<Rec>(g, mid, other, n, rest, &p2);
        if (passes_run) *passes_run = p1 + p2;
        return res;
    }
    const int np = (int)passes.size();
    PassList pl{};
    pl.n = np;
    for (int p = 0; p < np; ++p) {
        pl.field[p] = passes[p].field;
        pl.shift[p] = passes[p].shift;
    }
    const uint64_t n_tiles = (n + RS_TILE - 1) / RS_TILE;
    // hist: [np][256] counts, then [np][256] bin bases, then the ticket
    const size_t hist_bytes = (size_t)np * 256 * 8;
    g->hist.reserve(2 * hist_bytes + 256, s);
    unsigned long long *d_hist = g->hist.as<unsigned long long>();
    uint64_t *d_base = g->hist.as<uint64_t>() + (size_t)np * 256;
    unsigned int *d_ticket = reinterpret_cast<unsigned int *>(g->hist.as<unsigned char>() + 2 * hist_bytes);
    PDG_CUDA(cudaMemsetAsync(d_hist, 0, hist_bytes, s));
    unsigned hist_grid = (unsigned)std::min<uint64_t>((n + RS_THREADS - 1) / RS_THREADS, 148 * 8);
    radix_hist_kernel<Rec><<<hist_grid, RS_THREADS, 0, s>>>(a, n, pl, d_hist);
    PDG_LAUNCHED(g);
    std::vector<uint64_t> h_hist((size_t)np * 256), h_base((size_t)np * 256);
    PDG_CUDA(cudaMemcpyAsync(h_hist.data(), d_hist, hist_bytes, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    std::vector<int> run;
    for (int p = 0; p < np; ++p) {
        bool constant = false;
        uint64_t acc = 0;
        for (int d = 0; d < 256; ++d) {
            h_base[(size_t)p * 256 + d] = acc;
            acc += h_hist[(size_t)p * 256 + d];
            if (h_hist[(size_t)p * 256 + d] == n) constant = true;
        }
        if (!constant) run.push_back(p);
    }
    PDG_CUDA(cudaMemcpyAsync(d_base, h_base.data(), hist_bytes, cudaMemcpyHostToDevice, s));

    const size_t smem = sizeof(Rec) * RS_TILE + sizeof(uint32_t) * RS_WARPS * 256;
    const uint32_t which = sizeof(Rec) == 16 ? pdg_graph::ATTR_RADIX16 : pdg_graph::ATTR_RADIX8;
    if (!(g->attr_mask & which)) {
        PDG_CUDA(cudaFuncSetAttribute(onesweep_kernel<Rec>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        g->attr_mask |= which;
    }
    const size_t status_bytes = n_tiles * 256 * sizeof(unsigned long long);
    g->status.reserve(status_bytes, s);
    const bool timed = g->collect_sort_stats;
    if (timed && !g->ev0) {
        PDG_CUDA(cudaEventCreate(&g->ev0));
        PDG_CUDA(cudaEventCreate(&g->ev1));
    }
    Rec *src = a, *dst = b;
    if (timed && !run.empty()) PDG_CUDA(cudaEventRecord(g->ev0, s));
    for (int p : run) {
        PDG_CUDA(cudaMemsetAsync(g->status.p, 0, status_bytes, s));
        PDG_CUDA(cudaMemsetAsync(d_ticket, 0, sizeof(unsigned int), s));
        onesweep_kernel<Rec><<<(unsigned)n_tiles, RS_THREADS, smem, s>>>(
            src, dst, n, pl.field[p], pl.shift[p], d_base + (size_t)p * 256,
            g->status.as<unsigned long long>(), d_ticket);
        PDG_LAUNCHED(g);
        std::swap(src, dst);
    }
    if (timed && !run.empty()) {
        PDG_CUDA(cudaEventRecord(g->ev1, s));
        PDG_CUDA(cudaEventSynchronize(g->ev1));
        float ms = 0;
        cudaEventElapsedTime(&ms, g->ev0, g->ev1);
        g->stats.ms_sort_kernel_avg = ms / (float)run.size();   // includes the status memsets
        g->stats.sort_kernel_launches = run.size();
    }
    if (passes_run) *passes_run = run.size();
    return src;
}

EdgeRec *radix_sort_records(pdg_graph *g, EdgeRec *a, EdgeRec *b, uint64_t n,
                            const std::vector<PassDesc> &passes, uint64_t *passes_run) {
    return radix_sort_impl<EdgeRec>(g, a, b, n, passes, passes_run);
}

uint64_t *radix_sort_keys(pdg_graph *g, uint64_t *a, uint64_t *b, uint64_t n, int lo_bit, int hi_bit,
                          uint64_t *passes_run) {
    std::vector<PassDesc> passes;
    for (int sft = lo_bit; sft < hi_bit; sft += 8) passes.push_back(PassDesc{0, sft});
    return radix_sort_impl<uint64_t>(g, a, b, n, passes, passes_run);
}

}  // namespace pdg
