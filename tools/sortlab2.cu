// sortlab2.cu -- CTA-level partition kernel with PRECOMPUTED per-tile offsets (no look-back):
// how fast can one 8-byte-record partition pass go?  Variants by template parameters.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}
// records: (22-bit random code << 32) | index
__global__ void gen_kernel(uint64_t *r, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) r[i] = ((mix64(i * 0x9e3779b97f4a7c15ULL + 777) & 0x3FFFFF) << 32) | i;
}

template <int THREADS, int ITEMS, int BITS>
__global__ void __launch_bounds__(THREADS) tile_hist_kernel(const uint64_t *__restrict__ in, uint32_t n, int shift,
                                                            uint32_t *__restrict__ H) {
    constexpr int BINS = 1 << BITS, TILE = THREADS * ITEMS;
    __shared__ uint32_t cnt[BINS];
    for (int i = threadIdx.x; i < BINS; i += THREADS) cnt[i] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * TILE;
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const uint32_t i = base + j * THREADS + threadIdx.x;
        if (i < n) atomicAdd(&cnt[(uint32_t)(in[i] >> (32 + shift)) & (BINS - 1)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < BINS; i += THREADS) H[(size_t)blockIdx.x * BINS + i] = cnt[i];
}
template <int BINS>
__global__ void col_scan_kernel(const uint32_t *__restrict__ H, uint32_t n_tiles, uint32_t *__restrict__ O) {
    __shared__ uint32_t tot[BINS];
    const int d = threadIdx.x;
    uint32_t t = 0;
    for (uint32_t s = 0; s < n_tiles; ++s) t += H[(size_t)s * BINS + d];
    tot[d] = t;
    __syncthreads();
    uint32_t base = 0;
    for (int i = 0; i < d; ++i) base += tot[i];
    for (uint32_t s = 0; s < n_tiles; ++s) { O[(size_t)s * BINS + d] = base; base += H[(size_t)s * BINS + d]; }
}

// scatter with per-tile offsets.  OUT4: write only the low 32 bits (position column).
template <int THREADS, int ITEMS, int BITS, int MINB, bool OUT4, int RANK>
__global__ void __launch_bounds__(THREADS, MINB)
scatter_kernel(const uint64_t *__restrict__ in, uint32_t n, int shift, const uint32_t *__restrict__ O,
               uint64_t *__restrict__ out8, uint32_t *__restrict__ out4) {
    constexpr int BINS = 1 << BITS, TILE = THREADS * ITEMS, WARPS = THREADS / 32;
    extern __shared__ __align__(16) unsigned char smem[];
    uint64_t *stage = reinterpret_cast<uint64_t *>(smem);                       // [TILE]
    uint32_t *cnt = reinterpret_cast<uint32_t *>(smem + 8 * TILE);              // [WARPS][BINS] (x2 words when RANK==1)
    uint32_t *delta = cnt + WARPS * BINS * (RANK == 1 ? 2 : 1);                                       // [BINS] global offset - start in tile
    uint32_t *bstart = delta + BINS;                                            // [BINS + 1]
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t tile = blockIdx.x, base = tile * TILE;
    const uint32_t tile_n = min((uint32_t)TILE, n - base);
    for (int i = tid; i < WARPS * BINS * (RANK == 1 ? 2 : 1); i += THREADS) cnt[i] = 0;
    __syncthreads();
    uint64_t r[ITEMS];
    uint32_t dg[ITEMS];
    const uint32_t wbase = w * (ITEMS * 32);
    if (tile_n == TILE) {
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            r[j] = in[base + wbase + j * 32 + lane];
            dg[j] = ((uint32_t)(r[j] >> 32) >> shift) & (BINS - 1);
        }
    } else {
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t li = wbase + j * 32 + lane;
            r[j] = li < tile_n ? in[base + li] : 0;
            dg[j] = li < tile_n ? (((uint32_t)(r[j] >> 32) >> shift) & (BINS - 1)) : 0xFFFFFFFFu;
        }
    }
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t rank[ITEMS];
    uint32_t *wcnt = cnt + w * BINS * (RANK == 1 ? 2 : 1);
    if (RANK == 0) {
        uint32_t peers[ITEMS];
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) peers[j] = __match_any_sync(0xffffffffu, dg[j]);
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            uint32_t old = 0;
            if ((peers[j] & lt) == 0 && dg[j] != 0xFFFFFFFFu) old = atomicAdd(&wcnt[dg[j]], (uint32_t)__popc(peers[j]));
            rank[j] = old;
        }
#pragma unroll
        for (int j = 0; j < ITEMS; ++j)
            rank[j] = __shfl_sync(0xffffffffu, rank[j], __ffs(peers[j]) - 1) + __popc(peers[j] & lt);
    } else {
        // per (warp, bin) a pair {lane mask of this round, running count}: every lane ORs its bit in,
        // reads the pair back, and the lowest lane of each group folds the mask into the count
        uint2 *wt = reinterpret_cast<uint2 *>(wcnt);
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const bool ok = dg[j] != 0xFFFFFFFFu;
            if (ok) atomicOr(&wt[dg[j]].x, 1u << lane);
            __syncwarp();
            uint2 mc = make_uint2(0, 0);
            if (ok) mc = wt[dg[j]];
            rank[j] = mc.y + __popc(mc.x & lt);
            __syncwarp();
            if (ok && (mc.x & lt) == 0) wt[dg[j]] = make_uint2(0u, mc.y + __popc(mc.x));
            __syncwarp();
        }
    }
    __syncthreads();
    // cross-warp exclusive prefix per bin + tile count, then scan of the bins (warp 0.. via shuffles)
    for (int b = tid; b < BINS; b += THREADS) {
        uint32_t run = 0;
#pragma unroll
        for (int ww = 0; ww < WARPS; ++ww) {
            constexpr int ST = RANK == 1 ? 2 : 1;
            const uint32_t c = cnt[(ww * BINS + b) * ST + (ST - 1)];
            cnt[(ww * BINS + b) * ST + (ST - 1)] = run;
            run += c;
        }
        bstart[b] = run;   // count for now
    }
    __syncthreads();
    if (w == 0) {   // exclusive scan of BINS counts by one warp
        constexpr int PER = BINS / 32;
        uint32_t v[PER], s = 0;
#pragma unroll
        for (int i = 0; i < PER; ++i) { v[i] = bstart[lane * PER + i]; s += v[i]; }
        uint32_t inc = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
        uint32_t ex = inc - s;
        const uint32_t *Ot = O + (size_t)tile * BINS;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            bstart[lane * PER + i] = ex;
            delta[lane * PER + i] = Ot[lane * PER + i] - ex;
            ex += v[i];
        }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < ITEMS; ++j)
        if (dg[j] != 0xFFFFFFFFu) stage[bstart[dg[j]] + wcnt[dg[j] * (RANK == 1 ? 2 : 1) + (RANK == 1 ? 1 : 0)] + rank[j]] = r[j];
    __syncthreads();
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const uint32_t s = j * THREADS + tid;
        if (s < tile_n) {
            const uint64_t x = stage[s];
            const uint32_t d = ((uint32_t)(x >> 32) >> shift) & (BINS - 1);
            if (OUT4) out4[delta[d] + s] = (uint32_t)x; else out8[delta[d] + s] = x;
        }
    }
}

__global__ void verify_kernel(const uint64_t *__restrict__ out, uint32_t n, int shift, uint32_t mask, unsigned long long *bad) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0 || i >= n) return;
    const uint64_t a = out[i - 1], b = out[i];
    const uint64_t ka = ((uint64_t)(((uint32_t)(a >> 32) >> shift) & mask) << 32) | (uint32_t)a;
    const uint64_t kb = ((uint64_t)(((uint32_t)(b >> 32) >> shift) & mask) << 32) | (uint32_t)b;
    if (ka >= kb) atomicAdd(bad, 1ULL);
}

template <typename F> static float time_it(F f, int reps = 5) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    f();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int i = 0; i < reps; ++i) {
        CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    return best;
}

template <int THREADS, int ITEMS, int BITS, int MINB, bool OUT4, int RANK>
static void run(const uint64_t *in, uint32_t n, uint32_t *H, uint32_t *O, uint64_t *out, unsigned long long *bad) {
    constexpr int BINS = 1 << BITS, TILE = THREADS * ITEMS, WARPS = THREADS / 32;
    const int shift = 22 - BITS;
    const uint32_t n_tiles = (n + TILE - 1) / TILE;
    const size_t smem = 8 * TILE + 4 * (WARPS * BINS * (RANK == 1 ? 2 : 1) + 2 * BINS + 1);
    CK(cudaFuncSetAttribute(scatter_kernel<THREADS, ITEMS, BITS, MINB, OUT4, RANK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, scatter_kernel<THREADS, ITEMS, BITS, MINB, OUT4, RANK>, THREADS, smem));
    float th = time_it([&] { tile_hist_kernel<THREADS, ITEMS, BITS><<<n_tiles, THREADS>>>(in, n, shift, H); });
    float ts = time_it([&] { col_scan_kernel<BINS><<<1, BINS>>>(H, n_tiles, O); });
    float tp = time_it([&] { scatter_kernel<THREADS, ITEMS, BITS, MINB, OUT4, RANK><<<n_tiles, THREADS, smem>>>(in, n, shift, O, out, (uint32_t *)out); });
    unsigned long long hb = 0;
    if (!OUT4) {
        CK(cudaMemset(bad, 0, 8));
        verify_kernel<<<(n + 255) / 256, 256>>>(out, n, shift, BINS - 1, bad);
        CK(cudaMemcpy(&hb, bad, 8, cudaMemcpyDeviceToHost));
    }
    printf("rank=%d T=%d I=%d bits=%d minb=%d out4=%d occ=%d smem=%zu: tile_hist %.3f  scan %.3f  scatter %.3f ms (%.0f GB/s r+w)  bad=%llu\n",
           RANK, THREADS, ITEMS, BITS, MINB, (int)OUT4, occ, smem, th, ts, tp, n * (OUT4 ? 12.0 : 16.0) / tp / 1e6, hb);
}

int main(int argc, char **argv) {
    const uint32_t n = argc > 1 ? (uint32_t)strtoull(argv[1], 0, 10) : 100000000u;
    uint64_t *in, *out;
    uint32_t *H, *O;
    unsigned long long *bad;
    CK(cudaMalloc(&in, (size_t)n * 8)); CK(cudaMalloc(&out, (size_t)n * 8));
    CK(cudaMalloc(&H, (size_t)256 << 20)); CK(cudaMalloc(&O, (size_t)256 << 20)); CK(cudaMalloc(&bad, 8));
    gen_kernel<<<(n + 255) / 256, 256>>>(in, n);
    CK(cudaDeviceSynchronize());
    run<512, 8, 8, 3, false, 0>(in, n, H, O, out, bad);
    run<512, 8, 8, 3, false, 1>(in, n, H, O, out, bad);
    run<512, 8, 7, 3, false, 1>(in, n, H, O, out, bad);
    run<512, 8, 8, 2, false, 1>(in, n, H, O, out, bad);
    run<256, 8, 8, 6, false, 1>(in, n, H, O, out, bad);
    run<256, 16, 8, 3, false, 1>(in, n, H, O, out, bad);
    run<256, 16, 7, 3, false, 1>(in, n, H, O, out, bad);
    run<1024, 4, 8, 2, false, 1>(in, n, H, O, out, bad);
    run<512, 16, 8, 1, false, 1>(in, n, H, O, out, bad);
    run<512, 8, 8, 3, true, 1>(in, n, H, O, out, bad);
    return 0;
}
