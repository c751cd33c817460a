// noalloc.cu -- pdg_codes_checksum_static: the node checksum of pdg_checksum (export.cu: checksum_nodes_kernel) for
// callers that must not allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
//
// The existing path reserves a device accumulator per call (DevBuf::reserve -> cudaMalloc).  This variant keeps its
// scratch -- one 64-bit partial sum per CTA -- in a file-scope __device__ array of fixed size: NA_MAX_PARTIALS x 8 B =
// 1 MiB, which every GPU that loads the library holds until the process exits.  A CTA sums NA_CODES_PER_CTA codes, so
// the entry point takes at most PDG_STATIC_MAX_CODES = 2^27 codes (the node table of config 3 at k = 31 has 1.28e8) and
// refuses more.  It calls no cudaMalloc / cudaFree and no allocating library routine, and never synchronises: two
// kernel launches on the caller's stream, inputs and the output in the caller's device memory.
//
// Every call shares the array.  Nothing in it has to start at zero: a call's second kernel reads exactly the partials
// its first kernel wrote, so what the previous call left behind is never seen.  Calls that OVERLAP on the GPU
// (different streams, or different host threads without an order between their launches) race on it and must be
// serialised by the caller.
// No counterpart in the reference.
#include "common.cuh"
#include <mutex>

namespace pdg {

static constexpr int NA_THREADS = 256, NA_ITEMS = 4, NA_CODES_PER_CTA = NA_THREADS * NA_ITEMS;
static constexpr uint32_t NA_MAX_PARTIALS = 1u << 17;
static_assert((uint64_t)NA_MAX_PARTIALS * NA_CODES_PER_CTA == PDG_STATIC_MAX_CODES, "header and scratch size disagree");

static __device__ __align__(256) unsigned long long na_partials[NA_MAX_PARTIALS];

// partial[c] = sum over the CTA's codes of mix64(code + golden ratio), modulo 2^64 (order independent)
__global__ void __launch_bounds__(NA_THREADS) na_partial_kernel(const uint64_t *__restrict__ codes, uint64_t n,
                                                                unsigned long long *__restrict__ partial) {
    __shared__ unsigned long long wsum[NA_THREADS / 32];
    const uint64_t base = (uint64_t)blockIdx.x * NA_CODES_PER_CTA + threadIdx.x;
    unsigned long long h = 0;
#pragma unroll
    for (int j = 0; j < NA_ITEMS; ++j) {
        const uint64_t i = base + (uint64_t)j * NA_THREADS;
        if (i < n) h += mix64(codes[i] + 0x9e3779b97f4a7c15ULL);
    }
    for (int o = 16; o; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = h;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
#pragma unroll
        for (int w = 0; w < NA_THREADS / 32; ++w) t += wsum[w];
        partial[blockIdx.x] = t;
    }
}

// one CTA: *out = sum of partial[0, n_partials)  (0 when there are none)
__global__ void __launch_bounds__(1024) na_final_kernel(const unsigned long long *__restrict__ partial, uint32_t n_partials,
                                                        unsigned long long *__restrict__ out) {
    __shared__ unsigned long long wsum[32];
    unsigned long long h = 0;
    for (uint32_t i = threadIdx.x; i < n_partials; i += 1024) h += partial[i];
    for (int o = 16; o; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = h;
    __syncthreads();
    if (threadIdx.x < 32) {
        h = wsum[threadIdx.x];
        for (int o = 16; o; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
        if (threadIdx.x == 0) *out = h;
    }
}

// Device address of na_partials on `device` (which must be current).  In host code the array's NAME is only a
// host-side stand-in for the symbol, not a device address: it is resolved once per device and cached.
static constexpr int NA_MAX_DEVICES = 64;
static std::mutex na_mu;
static unsigned long long *na_addr[NA_MAX_DEVICES];

static unsigned long long *na_scratch(int device) {
    std::lock_guard<std::mutex> lock(na_mu);
    if (!na_addr[device]) {
        unsigned long long *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **)&p, na_partials);
        if (e != cudaSuccess || !p) {
            set_error("pdg_codes_checksum_static: cudaGetSymbolAddress on device %d -> %s", device, cudaGetErrorString(e));
            return nullptr;
        }
        na_addr[device] = p;
    }
    return na_addr[device];
}

}  // namespace pdg

using namespace pdg;

extern "C" int pdg_codes_checksum_static(int device, const uint64_t *d_codes, uint64_t n, uint64_t *d_sum,
                                         void *cuda_stream) {
    if (n > PDG_STATIC_MAX_CODES) {
        set_error("pdg_codes_checksum_static: %llu codes exceed the static scratch (at most %llu codes per call)",
                  (unsigned long long)n, (unsigned long long)PDG_STATIC_MAX_CODES);
        return PDG_E_CAPACITY;
    }
    if (!d_sum || (n && !d_codes)) {
        set_error("pdg_codes_checksum_static: null %s pointer", d_sum ? "input" : "output");
        return PDG_E_INVALID;
    }
    if (device < 0 || device >= NA_MAX_DEVICES) {
        set_error("pdg_codes_checksum_static: device %d out of range [0, %d)", device, NA_MAX_DEVICES);
        return PDG_E_INVALID;
    }
    int prev = -1;
    if (cudaGetDevice(&prev) != cudaSuccess || (prev != device && cudaSetDevice(device) != cudaSuccess)) {
        set_error("pdg_codes_checksum_static: cannot make device %d current: %s", device,
                  cudaGetErrorString(cudaGetLastError()));
        return PDG_E_CUDA;
    }
    int rc = PDG_OK;
    unsigned long long *scratch = na_scratch(device);
    if (!scratch) {
        rc = PDG_E_CUDA;
    } else {
        cudaStream_t s = (cudaStream_t)cuda_stream;
        const uint32_t n_partials = (uint32_t)((n + NA_CODES_PER_CTA - 1) / NA_CODES_PER_CTA);
        if (n_partials) na_partial_kernel<<<n_partials, NA_THREADS, 0, s>>>(d_codes, n, scratch);
        na_final_kernel<<<1, 1024, 0, s>>>(scratch, n_partials, (unsigned long long *)d_sum);
        const cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) {
            set_error("pdg_codes_checksum_static: launch failed: %s", cudaGetErrorString(e));
            rc = PDG_E_CUDA;
        }
    }
    if (prev != device) cudaSetDevice(prev);
    return rc;
}
