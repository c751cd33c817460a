// common.cuh -- shared declarations of libpdgpu (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>
#include "../../include/pdgpu.h"

namespace pdg {

void set_error(const char *fmt, ...);

struct CudaError { int code; };

#define PDG_CUDA(expr)                                                                   \
    do {                                                                                 \
        cudaError_t _e = (expr);                                                         \
        if (_e != cudaSuccess) {                                                         \
            pdg::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,                 \
                           cudaGetErrorString(_e));                                      \
            throw pdg::CudaError{_e == cudaErrorMemoryAllocation ? PDG_E_NOMEM           \
                                                                 : PDG_E_CUDA};          \
        }                                                                                \
    } while (0)

#define PDG_LAUNCH_CHECK() PDG_CUDA(cudaGetLastError())

// A growable device allocation.  Contents are preserved on growth when keep == true.
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    ~DevBuf() { release(); }
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    // ensure at least `bytes`; new space is zero filled when zero == true
    void reserve(size_t bytes, cudaStream_t s, bool keep = false, bool zero = false) {
        if (bytes <= cap) return;
        size_t ncap = keep ? (bytes + bytes / 2 + 256) : bytes;
        ncap = (ncap + 255) & ~size_t(255);
        void *np = nullptr;
        PDG_CUDA(cudaMalloc(&np, ncap));
        if (zero) PDG_CUDA(cudaMemsetAsync(np, 0, ncap, s));
        if (keep && p && cap) PDG_CUDA(cudaMemcpyAsync(np, p, cap, cudaMemcpyDeviceToDevice, s));
        if (p) {
            PDG_CUDA(cudaStreamSynchronize(s));
            cudaFree(p);
        }
        p = np;
        cap = ncap;
    }
    template <typename T> T *as() const { return reinterpret_cast<T *>(p); }
};

// one radix pass over a record field: digit = (field >> shift) & 0xFF
struct PassDesc { int field; int shift; };

// edge record while sorting: x = src id, y = tgt id, z = type id, w = value ^ 0x80000000
typedef uint4 EdgeRec;

static constexpr int kBitmapMaxBits = 30;   // presence-bitmap node path for codes < 2^30 (k <= 15)

}  // namespace pdg

#ifdef __CUDACC__
namespace pdg {
__device__ __forceinline__ uint64_t warp_incl_scan(uint64_t v) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint64_t t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

// block-wide exclusive scan of one value per thread (blockDim.x a multiple of 32, <= 1024);
// returns the exclusive prefix, *total = block sum (same value in every thread)
__device__ __forceinline__ uint64_t block_excl_scan(uint64_t v, uint64_t *total) {
    __shared__ uint64_t ws[33];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const uint64_t inc = warp_incl_scan(v);
    __syncthreads();   // a previous call may still be reading ws
    if (lane == 31) ws[w] = inc;
    __syncthreads();
    if (w == 0) {
        const uint64_t s = lane < nw ? ws[lane] : 0;
        const uint64_t si = warp_incl_scan(s);
        ws[lane] = si - s;            // exclusive base of warp `lane`
        if (lane == 31) ws[32] = si;  // block total
    }
    __syncthreads();
    if (total) *total = ws[32];
    return ws[w] + inc - v;
}

__device__ __forceinline__ uint64_t mix64(uint64_t x) {   // murmur3 finaliser
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}
}  // namespace pdg
#endif

struct pdg_graph {
    int device = 0;
    int k = 11;
    cudaStream_t stream = nullptr;

    // ---- genome store: residues packed 2 bits each, big endian inside 64-bit words ----
    pdg::DevBuf packed;            // uint64 words, zero beyond n_res
    uint64_t n_res = 0;
    std::vector<uint64_t> c_begin, c_len, c_edge_base;   // valid contigs (len > k), ascending begin
    std::vector<uint32_t> c_type;
    uint64_t n_kmers = 0;          // sum(len-k+1) over valid contigs
    uint64_t n_gen_edges = 0;      // sum(len-k)
    bool contigs_sorted_fresh = true;   // type ids strictly increasing with begin => fast path

    // ---- explicit (per-edge API) staging, host side until finalize ----
    std::vector<uint64_t> x_src, x_tgt, x_nodes;
    std::vector<uint32_t> x_type;
    std::vector<int32_t> x_val;

    // ---- finalized graph ----
    bool finalized = false;
    uint64_t n_nodes = 0, n_edges = 0;
    pdg::DevBuf node_codes;        // uint64[n_nodes], ascending
    pdg::DevBuf row_ptr, col_ptr;  // uint64[n_nodes+1]
    pdg::DevBuf csr_tgt, csr_type, csr_val;   // uint32[n_edges]; val biased (^0x80000000)
    pdg::DevBuf csc_src, csc_type, csc_val;
    // node lookup: presence bitmap + exclusive popcount prefix (bitmap path)
    bool have_bitmap = false;
    uint64_t bitmap_words = 0;     // 32-bit words
    pdg::DevBuf bitmap, bitmap_prefix;

    // ---- scratch ----
    pdg::DevBuf d_cbegin, d_cend, d_cebase, d_ctype;     // device copy of the contig table
    pdg::DevBuf recA, recB, scratch0, scratch1, scratch2, status, hist;
    pdg::DevBuf q_a, q_b, q_cnt, q_off, q_hits, small;
    uint64_t q_n = 0, q_total = 0;

    pdg_build_stats stats{};

    // ---- multi GPU ----
    void *nccl_comm = nullptr;
    int rank = 0, world = 1;
};

namespace pdg {

// ---- scan.cu ----
// out[i] = sum(in[0..i)) ; out has n+1 entries (out[n] = total).  in: uint32, out: uint64.
void exclusive_scan_u32_u64(const uint32_t *d_in, uint64_t *d_out, uint64_t n, pdg_graph *g);
void exclusive_scan_u32_u32(const uint32_t *d_in, uint32_t *d_out, uint64_t n, pdg_graph *g);

// ---- radix.cu ----
// LSD onesweep radix sort.  Sorts n records from `a` using `b` as the other buffer, running the
// given passes in order; passes whose digit is constant over the input are skipped.  Returns the
// buffer (a or b) that holds the result.  Stable.
EdgeRec *radix_sort_records(pdg_graph *g, EdgeRec *a, EdgeRec *b, uint64_t n,
                            const std::vector<PassDesc> &passes, uint64_t *passes_run);
uint64_t *radix_sort_keys(pdg_graph *g, uint64_t *a, uint64_t *b, uint64_t n, int key_bits,
                          uint64_t *passes_run);

// ---- ingest.cu ----
int ingest_fasta(pdg_graph *g, const uint8_t *d_bytes, uint64_t n_bytes, uint64_t *h_header_table,
                 uint64_t cap_headers, pdg_ingest_info *info);

// ---- build.cu ----
void upload_contigs(pdg_graph *g);
int finalize(pdg_graph *g, int flags);
int kmer_stats(pdg_graph *g, const uint64_t *h_begin, const uint64_t *h_len, uint64_t n,
               uint64_t *n_kmers, uint64_t *n_unique);
int emit_codes(pdg_graph *g, uint64_t begin, uint64_t len, uint64_t *d_out);
int checksum(pdg_graph *g, uint64_t *nodes_sum, uint64_t *edges_sum);

// ---- query.cu ----
int query_run(pdg_graph *g, const uint64_t *h_a, const uint64_t *h_b, uint64_t n_q, int64_t max_len,
              bool connected_only, uint64_t *total);
int query_walk(pdg_graph *g, uint64_t src_code, const pdg_hit *hit, uint32_t *h_ids, int32_t *status);

// ---- comm.cu ----
// replaces g->node_codes (this rank's sorted unique codes) by the global table (all ranks)
int comm_global_nodes(pdg_graph *g, int key_bits);

}  // namespace pdg
