// export.cu -- reading a finalized graph back: k-mer statistics, codes / residues of a range,
// the labelled edge list, dense node ids and order-independent checksums.
//   prairiedog/kmers.py:48-49,134-150   Kmers.__len__ / unique_kmers        -> kmer_stats
//   prairiedog/lemon_graph.py:133-149   LGGraph.nodes / edges               -> export_edges
#include "gindex.cuh"
#include <algorithm>

namespace pdg {

// ---- genome part edges, in index order -------------------------------------------------------
// occurrence s (position p, row r) is an out-edge unless it is the last k-mer of its contig
__global__ void g_edge_flags_kernel(const uint32_t *__restrict__ pos, uint64_t n, int k, ContigTable ct,
                                    uint32_t *__restrict__ flags) {
    const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const uint64_t p = pos[s];
    flags[s] = (p + k < ct.end[contig_of(ct, p)]) ? 1u : 0u;
}

// row of occurrence s: last r with rowptr[r] <= s
__device__ __forceinline__ uint64_t row_of(const uint32_t *rowptr, uint64_t n_rows, uint64_t s) {
    uint64_t lo = 0, hi = n_rows;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (rowptr[mid] <= s) lo = mid + 1; else hi = mid;
    }
    return lo - 1;
}

// MODE 0: write edges at off[s];  MODE 1: add the mixed hash of every edge into *sum
template <int MODE>
__global__ void g_edges_kernel(const unsigned long long *__restrict__ packed, const uint32_t *__restrict__ pos,
                               const uint32_t *__restrict__ rowptr, const uint64_t *__restrict__ codes,
                               uint64_t n_rows, uint64_t n, int k, ContigTable ct,
                               const uint32_t *__restrict__ flags, const uint32_t *__restrict__ off,
                               uint64_t *__restrict__ o_src, uint64_t *__restrict__ o_tgt,
                               uint32_t *__restrict__ o_type, int32_t *__restrict__ o_val,
                               unsigned long long *sum) {
    const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long h = 0;
    if (s < n) {
        const uint64_t p = pos[s];
        const uint64_t c = contig_of(ct, p);
        if (p + k < ct.end[c]) {
            const uint64_t src = codes[row_of(rowptr, n_rows, s)];   // the row the index filed p under
            const uint64_t tgt = kmer_at(packed, p + 1, k);
            const uint32_t ty = ct.type[c];
            const int32_t v = (int32_t)(p - ct.begin[c]);
            if (MODE == 0) {
                const uint32_t o = off[s];
                if (o_src) o_src[o] = src;
                if (o_tgt) o_tgt[o] = tgt;
                if (o_type) o_type[o] = ty;
                if (o_val) o_val[o] = v;
            } else {
                h = mix64(src + 0x9e3779b97f4a7c15ULL);
                h = mix64(h ^ tgt);
                h = mix64(h ^ (((uint64_t)ty << 32) | (uint32_t)v));
                if (kmer_at(packed, p, k) != src) h = ~h;   // misfiled occurrence poisons the sum
            }
        }
    }
    if (MODE == 1) {
        for (int o = 16; o; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
        if ((threadIdx.x & 31) == 0 && h) atomicAdd(sum, h);
    }
}

// ---- explicit part ---------------------------------------------------------------------------
__global__ void x_edges_kernel(const uint64_t *__restrict__ ptr, uint64_t n_nodes, uint64_t n_edges,
                               const uint32_t *__restrict__ other, const uint32_t *__restrict__ type,
                               const uint32_t *__restrict__ val, const uint64_t *__restrict__ codes, int src_is_row,
                               uint64_t *__restrict__ o_src, uint64_t *__restrict__ o_tgt,
                               uint32_t *__restrict__ o_type, int32_t *__restrict__ o_val,
                               unsigned long long *sum) {
    const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long h = 0;
    if (e < n_edges) {
        uint64_t lo = 0, hi = n_nodes;   // last row with ptr[row] <= e
        while (lo < hi) {
            const uint64_t mid = (lo + hi) >> 1;
            if (ptr[mid + 1] <= e) lo = mid + 1; else hi = mid;
        }
        const uint64_t rowc = codes[lo], othc = codes[other[e]];
        const uint64_t s = src_is_row ? rowc : othc, t = src_is_row ? othc : rowc;
        const int32_t v = (int32_t)(val[e] ^ 0x80000000u);
        if (sum) {
            h = mix64(s + 0x9e3779b97f4a7c15ULL);
            h = mix64(h ^ t);
            h = mix64(h ^ (((uint64_t)type[e] << 32) | (uint32_t)v));
        } else {
            if (o_src) o_src[e] = s;
            if (o_tgt) o_tgt[e] = t;
            if (o_type) o_type[e] = type[e];
            if (o_val) o_val[e] = v;
        }
    }
    if (sum) {
        for (int o = 16; o; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
        if ((threadIdx.x & 31) == 0 && h) atomicAdd(sum, h);
    }
}

__global__ void checksum_nodes_kernel(const uint64_t *__restrict__ codes, uint64_t n, unsigned long long *sum) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long h = i < n ? mix64(codes[i] + 0x9e3779b97f4a7c15ULL) : 0ULL;
    for (int o = 16; o; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
    if ((threadIdx.x & 31) == 0 && h) atomicAdd(sum, h);
}

// every labelled edge of the graph as device columns (src code, tgt code, type id, value): the genome part in index
// order, then the explicit part in CSR order.  Shared by pdg_export_edges and the filter (filter.cu).
int materialise_edges(pdg_graph *g, DevBuf &d_src, DevBuf &d_tgt, DevBuf &d_type, DevBuf &d_val) {
    cudaStream_t s = g->stream;
    const uint64_t ne = g->n_edges, ng = g->g_edges, nx = g->x_n_edges, nk = g->g_kmers;
    if (ne == 0) return PDG_OK;
    d_src.reserve(ne * 8, s); d_tgt.reserve(ne * 8, s); d_type.reserve(ne * 4, s); d_val.reserve(ne * 4, s);
    if (ng) {
        g->scratch0.reserve(nk * 4, s);
        g->scratch1.reserve((nk + 1) * 4, s);
        const ContigTable ct = g_table(g);
        g_edge_flags_kernel<<<grid_for(nk, 256), 256, 0, s>>>(g->g_pos.as<uint32_t>(), nk, g->k, ct,
                                                             g->scratch0.as<uint32_t>());
        PDG_LAUNCHED(g);
        exclusive_scan_u32_u32(g->scratch0.as<uint32_t>(), g->scratch1.as<uint32_t>(), nk, g);
        g_edges_kernel<0><<<grid_for(nk, 256), 256, 0, s>>>(
            g->packed.as<unsigned long long>(), g->g_pos.as<uint32_t>(), g->g_rowptr.as<uint32_t>(),
            g->g_codes.as<uint64_t>(), g->g_nodes, nk, g->k, ct, g->scratch0.as<uint32_t>(),
            g->scratch1.as<uint32_t>(), d_src.as<uint64_t>(), d_tgt.as<uint64_t>(), d_type.as<uint32_t>(),
            d_val.as<int32_t>(), nullptr);
        PDG_LAUNCHED(g);
    }
    if (nx) {
        x_edges_kernel<<<grid_for(nx, 256), 256, 0, s>>>(
            g->row_ptr.as<uint64_t>(), g->x_n_nodes, nx, g->csr_tgt.as<uint32_t>(), g->csr_type.as<uint32_t>(),
            g->csr_val.as<uint32_t>(), g->x_codes.as<uint64_t>(), 1, d_src.as<uint64_t>() + ng,
            d_tgt.as<uint64_t>() + ng, d_type.as<uint32_t>() + ng, d_val.as<int32_t>() + ng, nullptr);
        PDG_LAUNCHED(g);
    }
    return PDG_OK;
}

int export_edges(pdg_graph *g, uint64_t *h_src, uint64_t *h_tgt, uint32_t *h_type, int32_t *h_val) {
    cudaStream_t s = g->stream;
    const uint64_t ne = g->n_edges;
    if (ne == 0) return PDG_OK;
    DevBuf d_src, d_tgt, d_type, d_val;
    const int rc = materialise_edges(g, d_src, d_tgt, d_type, d_val);
    if (rc != PDG_OK) return rc;
    if (h_src) PDG_CUDA(cudaMemcpyAsync(h_src, d_src.p, ne * 8, cudaMemcpyDeviceToHost, s));
    if (h_tgt) PDG_CUDA(cudaMemcpyAsync(h_tgt, d_tgt.p, ne * 8, cudaMemcpyDeviceToHost, s));
    if (h_type) PDG_CUDA(cudaMemcpyAsync(h_type, d_type.p, ne * 4, cudaMemcpyDeviceToHost, s));
    if (h_val) PDG_CUDA(cudaMemcpyAsync(h_val, d_val.p, ne * 4, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    return PDG_OK;
}

int checksum(pdg_graph *g, uint64_t *nodes_sum, uint64_t *edges_sum) {
    cudaStream_t s = g->stream;
    if (g->u_lazy) {   // multi-rank: collective on the first call after pdg_finalize
        const int rc = comm_materialize_nodes(g);
        if (rc != PDG_OK) return rc;
    }
    DevBuf d;
    d.reserve(32, s);
    PDG_CUDA(cudaMemsetAsync(d.p, 0, 32, s));
    unsigned long long *sums = d.as<unsigned long long>();
    const uint64_t n_local_nodes = g->world > 1 ? g->n_nodes : g->n_nodes;
    if (n_local_nodes) {
        checksum_nodes_kernel<<<grid_for(n_local_nodes, 256), 256, 0, s>>>(g->u_codes, n_local_nodes, sums);
        PDG_LAUNCHED(g);
    }
    if (g->g_kmers) {
        g_edges_kernel<1><<<grid_for(g->g_kmers, 256), 256, 0, s>>>(
            g->packed.as<unsigned long long>(), g->g_pos.as<uint32_t>(), g->g_rowptr.as<uint32_t>(),
            g->g_codes.as<uint64_t>(), g->g_nodes, g->g_kmers, g->k, g_table(g), nullptr, nullptr, nullptr,
            nullptr, nullptr, nullptr, sums + 1);
        PDG_LAUNCHED(g);
    }
    if (g->x_n_edges) {
        x_edges_kernel<<<grid_for(g->x_n_edges, 256), 256, 0, s>>>(
            g->row_ptr.as<uint64_t>(), g->x_n_nodes, g->x_n_edges, g->csr_tgt.as<uint32_t>(),
            g->csr_type.as<uint32_t>(), g->csr_val.as<uint32_t>(), g->x_codes.as<uint64_t>(), 1, nullptr, nullptr,
            nullptr, nullptr, sums + 2);
        PDG_LAUNCHED(g);
        x_edges_kernel<<<grid_for(g->x_n_edges, 256), 256, 0, s>>>(
            g->col_ptr.as<uint64_t>(), g->x_n_nodes, g->x_n_edges, g->csc_src.as<uint32_t>(),
            g->csc_type.as<uint32_t>(), g->csc_val.as<uint32_t>(), g->x_codes.as<uint64_t>(), 0, nullptr, nullptr,
            nullptr, nullptr, sums + 3);
        PDG_LAUNCHED(g);
    }
    unsigned long long h[4];
    PDG_CUDA(cudaMemcpyAsync(h, sums, 32, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    if (h[2] != h[3]) {
        set_error("pdg_checksum: CSR and CSC of the explicit part disagree (%llx vs %llx)", h[2], h[3]);
        return PDG_E_INVALID;
    }
    *nodes_sum = h[0];
    *edges_sum = h[1] + h[2];
    return PDG_OK;
}

}  // namespace pdg
