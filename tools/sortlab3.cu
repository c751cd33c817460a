// sortlab3.cu -- per-SM throughput of the primitives a partition kernel leans on.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); return 1; } } while (0)

__device__ __forceinline__ uint32_t hash32(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

enum { OP_MATCH, OP_BALLOT8, OP_ATOMS_RET, OP_ATOMS_NORET, OP_ATOMS_RET_LEADER, OP_SHFL, OP_STS64_SCATTER, OP_LDS32_RANDOM, OP_LDS64_LINEAR, OP_POPC, OP_COUNT };
const char *NAMES[] = {"match_any(8-bit)", "8x ballot match", "ATOMS ret (all lanes, warp-private 256 bins)", "ATOMS noret", "ATOMS ret leader-only (1 lane)", "shfl idx", "STS.64 scattered", "LDS.32 random", "LDS.64 linear", "popc+alu"};

template <int OP>
__global__ void __launch_bounds__(512) prim_kernel(uint32_t iters, uint32_t *sink) {
    __shared__ uint32_t tab[16 * 256];
    __shared__ __align__(16) uint64_t st[4096];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    for (int i = tid; i < 16 * 256; i += 512) tab[i] = 0;
    for (int i = tid; i < 4096; i += 512) st[i] = i;
    __syncthreads();
    uint32_t x = hash32(tid + blockIdx.x * 977u), acc = 0;
    uint32_t *wt = tab + w * 256;
    for (uint32_t it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            x = x * 1664525u + 1013904223u;
            const uint32_t d = (x >> 13) & 255u;
            if (OP == OP_MATCH) acc += __match_any_sync(0xffffffffu, d);
            else if (OP == OP_BALLOT8) {
                uint32_t m = 0xffffffffu;
#pragma unroll
                for (int b = 0; b < 8; ++b) { const uint32_t v = __ballot_sync(0xffffffffu, (d >> b) & 1u); m &= ((d >> b) & 1u) ? v : ~v; }
                acc += m;
            } else if (OP == OP_ATOMS_RET) acc += atomicAdd(&wt[d], 1u);
            else if (OP == OP_ATOMS_NORET) atomicAdd(&wt[d], 1u);
            else if (OP == OP_ATOMS_RET_LEADER) { if (lane == (int)(d & 31u)) acc += atomicAdd(&wt[d], 1u); }
            else if (OP == OP_SHFL) acc += __shfl_sync(0xffffffffu, x, d & 31u);
            else if (OP == OP_STS64_SCATTER) st[(x >> 9) & 4095u] = ((uint64_t)x << 32) | d;
            else if (OP == OP_LDS32_RANDOM) acc += tab[(x >> 9) & 4095u];
            else if (OP == OP_LDS64_LINEAR) acc += (uint32_t)st[(tid + u * 512) & 4095u] + d;
            else if (OP == OP_POPC) acc += __popc(x & d) + (x >> 3);
        }
    }
    if (acc == 0x12345678u) sink[0] = acc + (uint32_t)st[lane] + tab[lane];
}

template <int OP> static int run(uint32_t *sink) {
    const uint32_t iters = 2000;
    const int grid = 148 * 3;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    prim_kernel<OP><<<grid, 512>>>(10, sink);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    prim_kernel<OP><<<grid, 512>>>(iters, sink);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    const double ops = (double)grid * 512 * iters * 8;
    // assume ~1.9 GHz for the cycles estimate
    printf("%-48s %8.3f ms  %7.1f G lane-ops/s  %.3f lane-ops/cycle/SM (at 1.9 GHz)\n", NAMES[OP], ms, ops / ms / 1e6, ops / (ms * 1e-3) / 148 / 1.9e9);
    return 0;
}

int main() {
    uint32_t *sink;
    CK(cudaMalloc(&sink, 64));
    run<OP_POPC>(sink); run<OP_MATCH>(sink); run<OP_BALLOT8>(sink); run<OP_ATOMS_RET>(sink); run<OP_ATOMS_NORET>(sink);
    run<OP_ATOMS_RET_LEADER>(sink); run<OP_SHFL>(sink); run<OP_STS64_SCATTER>(sink); run<OP_LDS32_RANDOM>(sink); run<OP_LDS64_LINEAR>(sink);
    return 0;
}
