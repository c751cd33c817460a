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
    uint64_t launches = 0;         // kernels launched since pdg_create / pdg_clear

    // ---- genome store: residues packed 2 bits each, big endian inside 64-bit words ----
    pdg::DevBuf packed;            // uint64 words, zero beyond n_res
    uint64_t n_res = 0;
    // every declared contig with len > k, ascending begin
    std::vector<uint64_t> c_begin, c_len;
    std::vector<uint32_t> c_type;
    uint64_t n_kmers = 0;          // sum(len-k+1) over declared contigs

    // ---- explicit (per-edge API) staging, host side until finalize ----
    std::vector<uint64_t> x_src, x_tgt, x_nodes;
    std::vector<uint32_t> x_type;
    std::vector<int32_t> x_val;

    bool finalized = false;
    uint64_t n_nodes = 0, n_edges = 0;   // whole graph (n_nodes: global when world > 1)

    // ---- G: genome part = occurrence index over the contigs that stayed in G ----
    std::vector<uint64_t> g_begin, g_len;    // host copy (ascending begin)
    std::vector<uint32_t> g_type;
    uint64_t g_kmers = 0, g_edges = 0, g_nodes = 0;
    pdg::DevBuf d_gbegin, d_gend, d_gkbase, d_gtype;   // device contig table (kbase: k-mers before)
    pdg::DevBuf g_codes;           // uint64[g_nodes] ascending = the rows
    pdg::DevBuf g_rowptr;          // uint32[g_nodes+1]
    pdg::DevBuf g_pos;             // uint32[g_kmers] residue positions, by (code, position)
    bool have_bitmap = false;      // presence bitmap + exclusive popcount prefix over g_codes
    uint64_t bitmap_words = 0;     // 32-bit words
    pdg::DevBuf bitmap, bitmap_prefix;

    // ---- X: explicit part (per-edge API content and demoted contigs) ----
    uint64_t x_n_nodes = 0, x_n_edges = 0;
    pdg::DevBuf x_codes;           // uint64[x_n_nodes] ascending
    pdg::DevBuf row_ptr, col_ptr;  // uint64[x_n_nodes+1]
    pdg::DevBuf csr_tgt, csr_type, csr_val;   // uint32[x_n_edges]; val biased (^0x80000000)
    pdg::DevBuf csc_src, csc_type, csc_val;

    // ---- node table of the whole graph (G union X; all ranks when world > 1) ----
    pdg::DevBuf u_codes_buf;
    const uint64_t *u_codes = nullptr;   // points at g_codes, x_codes or u_codes_buf

    // ---- scratch ----
    pdg::DevBuf m3_d1;                    // msd3.cu: low byte of the code of every pass-2 output record (pass 3 counts from it)
    bool m3_use_d1 = false;               // ... used by the staged groups since the last msd3_reserve (pipelined builds)
    pdg::DevBuf m3_tile_info, m3_gbase;   // msd3.cu: per-tile contig info; per-group cell bases + record offsets
    struct M3Group { uint64_t pos0, pos1, rec_off; };
    pdg::DevBuf ml_flag;                 // msdlong.cu: overflow flag + list of the buckets too large for a warp
    bool m3_launched = false;            // msd3_finish_launch ran for the staged groups
    std::vector<M3Group> m3_groups;      // groups staged since msd3_reserve
    uint32_t m3_max_groups = 0;
    bool m3_timed = false;
    // cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per DEVICE: remembered per graph (a graph lives on one
    // device), never in a process-wide static -- a second Engine on another GPU has to set it again
    uint32_t attr_mask = 0;
    enum { ATTR_MSD3 = 1, ATTR_MSDLONG = 2, ATTR_RADIX8 = 4, ATTR_RADIX16 = 8, ATTR_KSORT = 16 /* + 4 template bits */ };
    uint64_t m3_pos_lo = 0;
    pdg::DevBuf recA, recB, scratch0, scratch1, scratch2, status, hist;
    pdg::DevBuf q_a, q_b, q_cnt, q_cnt_g, q_off, q_hits, small;
    uint64_t q_n = 0, q_total = 0;
    bool q_connected = false;            // the last query call was pdg_query_connected
    pdg::DevBuf qs_len, qs_off, qs_bytes;   // strings of the last query's hits (query.cu)
    uint64_t qs_total = 0;
    bool qs_ready = false;
    // Graph.filter() working set (filter.cu), alive between pdg_filter_degrees and pdg_filter_fetch
    pdg::DevBuf flt_src, flt_tgt, flt_type, flt_val, flt_uid, flt_vid, flt_adj, flt_deg;
    pdg::DevBuf flt_osrc, flt_otgt, flt_otype, flt_oval;
    uint64_t flt_pairs = 0, flt_kept_edges = 0;
    bool flt_ready = false, flt_selected = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::vector<cudaEvent_t> k_events;   // per-kernel timing of msd3.cu / ingest_batch.cu
    bool collect_sort_stats = false;   // radix.cu times its pass kernels into stats

    pdg_build_stats stats{};

    // ---- batched ingest (ingest_batch.cu) ----
    pdg::DevBuf stage, fb_files, fb_summ, fb_prefix, fb_totals;
    cudaStream_t copy_stream = nullptr, aux_stream = nullptr;
    // pipelined build: contig table assumed at ingest time (device + host copy), valid until anything else touches the graph
    pdg::DevBuf pre_ct;
    std::vector<uint64_t> pre_begin, pre_len;
    bool pre_valid = false;
    cudaEvent_t fb_ev_ready = nullptr;
    std::vector<cudaEvent_t> fb_events, i_events;
    float ms_ingest[3] = {0, 0, 0};
    uint64_t ingest_bytes = 0, ingest_launches = 0;

    // ---- multi GPU ----
    pdg::DevBuf c_counts, c_all_counts, c_recv, c_own, c_tmp, c_all, c_all2;   // comm.cu scratch
    std::vector<uint64_t> c_own_counts;   // size of every rank's owned node set
    bool u_lazy = false;                  // multi-rank: the global node table is not gathered yet (comm_materialize_nodes)
    void *nccl_comm = nullptr;
    int rank = 0, world = 1;
};

#define PDG_LAUNCHED(g) do { ++(g)->launches; PDG_LAUNCH_CHECK(); } while (0)

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
// sorts on key bits [lo_bit, hi_bit)
uint64_t *radix_sort_keys(pdg_graph *g, uint64_t *a, uint64_t *b, uint64_t n, int lo_bit, int hi_bit,
                          uint64_t *passes_run);

// ---- ingest.cu ----
int ingest_fasta(pdg_graph *g, const uint8_t *d_bytes, uint64_t n_bytes, uint64_t *h_header_table,
                 uint64_t cap_headers, pdg_ingest_info *info);

// ---- ingest_batch.cu ----
int ingest_fasta_batch(pdg_graph *g, const uint8_t *const *ptrs, const uint64_t *n_bytes, uint32_t n_files,
                       int flags, uint64_t *h_table, uint64_t cap_headers, pdg_ingest_info *infos);

// ---- build.cu ----
int finalize(pdg_graph *g, int flags);
int kmer_stats(pdg_graph *g, const uint64_t *h_begin, const uint64_t *h_len, uint64_t n,
               uint64_t *n_kmers, uint64_t *n_unique);
int emit_codes(pdg_graph *g, uint64_t begin, uint64_t len, uint64_t *d_out);
int fetch_residues(pdg_graph *g, uint64_t begin, uint64_t len, uint8_t *h_ascii);
int checksum(pdg_graph *g, uint64_t *nodes_sum, uint64_t *edges_sum);
int export_edges(pdg_graph *g, uint64_t *h_src, uint64_t *h_tgt, uint32_t *h_type, int32_t *h_val);
int node_ids(pdg_graph *g, const uint64_t *h_codes, uint64_t n, uint64_t *h_ids);
// sorted unique codes of `keys` (n of them, destroyed) -> out (reserved here); returns the count
uint64_t sort_unique_keys(pdg_graph *g, uint64_t *keys, uint64_t *tmp, uint64_t n, int key_bits, DevBuf &out);

// ---- query.cu ----
int query_run(pdg_graph *g, const uint64_t *h_a, const uint64_t *h_b, uint64_t n_q, int64_t max_len,
              bool connected_only, uint64_t *total);
int query_walk(pdg_graph *g, uint64_t src_code, const pdg_hit *hit, uint64_t *h_codes, int32_t *status);
int query_strings_size(pdg_graph *g, uint64_t *total_bytes);
int query_strings_fetch(pdg_graph *g, uint64_t *h_offsets, uint8_t *h_bytes);

// ---- filter.cu ----
int filter_degrees(pdg_graph *g, uint32_t *max_degree);
int filter_select(pdg_graph *g, uint32_t bottom_degrees, uint64_t *n_kept_nodes, uint64_t *n_kept_edges);
int filter_fetch(pdg_graph *g, uint64_t *h_src, uint64_t *h_tgt, uint32_t *h_type, int32_t *h_val);

// ---- comm.cu ----
// in: sorted unique codes of this rank (d_local, n_local); out: g->u_codes_buf holds the global
// sorted unique table of all ranks, *n_global its size
int comm_global_nodes(pdg_graph *g, const uint64_t *d_local, uint64_t n_local, int key_bits, uint64_t *n_global);
// collective: gathers the global sorted node table into g->u_codes_buf if it is still distributed
int comm_materialize_nodes(pdg_graph *g);
void comm_destroy(pdg_graph *g);

}  // namespace pdg
