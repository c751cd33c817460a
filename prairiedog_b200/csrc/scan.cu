// scan.cu -- device-wide exclusive prefix sums used for bitmap ranks, run-length compaction
// offsets and query hit offsets.  Three launches: tile sums, one-CTA scan of tile sums, apply.
#include "common.cuh"

namespace pdg {

static constexpr int SCAN_THREADS = 256;
static constexpr int SCAN_ITEMS = 16;
static constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_sums(const uint32_t *__restrict__ in,
                                                               uint64_t *__restrict__ tile_sums,
                                                               uint64_t n) {
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE;
    uint64_t s = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; ++j) {
        uint64_t i = base + (uint64_t)j * SCAN_THREADS + threadIdx.x;
        if (i < n) s += in[i];
    }
    __shared__ uint64_t red[SCAN_THREADS / 32];
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t t = 0;
        for (int i = 0; i < SCAN_THREADS / 32; ++i) t += red[i];
        tile_sums[blockIdx.x] = t;
    }
}

// one CTA: in-place exclusive scan of tile sums; writes the grand total to tile_sums[n_tiles]
__global__ void __launch_bounds__(1024) scan_tile_offsets(uint64_t *tile_sums, uint64_t n_tiles) {
    uint64_t carry = 0;   // uniform across the block
    for (uint64_t base = 0; base < n_tiles; base += 1024) {
        const uint64_t i = base + threadIdx.x;
        const uint64_t v = i < n_tiles ? tile_sums[i] : 0;
        uint64_t tot;
        const uint64_t excl = block_excl_scan(v, &tot);
        if (i < n_tiles) tile_sums[i] = carry + excl;
        carry += tot;
    }
    if (threadIdx.x == 0) tile_sums[n_tiles] = carry;
}

template <typename OutT>
__global__ void __launch_bounds__(SCAN_THREADS) scan_apply(const uint32_t *__restrict__ in,
                                                           OutT *__restrict__ out,
                                                           const uint64_t *__restrict__ tile_offs,
                                                           uint64_t n, uint64_t n_tiles) {
    // thread t owns items [t*ITEMS, (t+1)*ITEMS) of the tile (blocked) for a sequential local scan
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE + (uint64_t)threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    uint64_t s = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; ++j) {
        uint64_t i = base + j;
        v[j] = i < n ? in[i] : 0u;
        s += v[j];
    }
    uint64_t excl = block_excl_scan(s, nullptr) + tile_offs[blockIdx.x];
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; ++j) {
        uint64_t i = base + j;
        if (i < n) out[i] = (OutT)excl;
        excl += v[j];
    }
    if (blockIdx.x == n_tiles - 1 && threadIdx.x == 0) out[n] = (OutT)tile_offs[n_tiles];
}

template <typename OutT>
static void scan_impl(const uint32_t *d_in, OutT *d_out, uint64_t n, pdg_graph *g) {
    cudaStream_t s = g->stream;
    if (n == 0) {
        PDG_CUDA(cudaMemsetAsync(d_out, 0, sizeof(OutT), s));
        return;
    }
    uint64_t n_tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    g->scratch2.reserve((n_tiles + 1) * sizeof(uint64_t), s);
    uint64_t *tiles = g->scratch2.as<uint64_t>();
    scan_tile_sums<<<(unsigned)n_tiles, SCAN_THREADS, 0, s>>>(d_in, tiles, n);
    PDG_LAUNCHED(g);
    scan_tile_offsets<<<1, 1024, 0, s>>>(tiles, n_tiles);
    PDG_LAUNCHED(g);
    scan_apply<OutT><<<(unsigned)n_tiles, SCAN_THREADS, 0, s>>>(d_in, d_out, tiles, n, n_tiles);
    PDG_LAUNCHED(g);
}

void exclusive_scan_u32_u64(const uint32_t *d_in, uint64_t *d_out, uint64_t n, pdg_graph *g) {
    scan_impl<uint64_t>(d_in, d_out, n, g);
}
void exclusive_scan_u32_u32(const uint32_t *d_in, uint32_t *d_out, uint64_t n, pdg_graph *g) {
    scan_impl<uint32_t>(d_in, d_out, n, g);
}

}  // namespace pdg
