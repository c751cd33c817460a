// stats.cu -- Kmers-level bulk calls on the genome store.
//   prairiedog/kmers.py:48-49,134-150   Kmers.__len__ / unique_kmers   -> kmer_stats
//   prairiedog/kmers.py:96-124          Kmers.next (bulk)              -> emit_codes
//   prairiedog/kmers.py:51-77           Kmers.sequences                -> fetch_residues
#include "gindex.cuh"
#include <algorithm>

namespace pdg {

__global__ void count_set_bits_kernel(const uint32_t *__restrict__ bm, uint64_t n_words, unsigned long long *sum) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long c = i < n_words ? __popc(bm[i]) : 0;
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(sum, c);
}
__global__ void count_heads_kernel(const uint64_t *__restrict__ keys, uint64_t n, unsigned long long *sum) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long c = (i < n && (i == 0 || keys[i] != keys[i - 1])) ? 1 : 0;
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(sum, c);
}
__global__ void decode_residues_kernel(const unsigned long long *__restrict__ packed, uint64_t begin, uint64_t len,
                                       uint8_t *__restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= len) return;
    out[i] = "ATCG"[base_at(packed, begin + i)];   // A=0 T=1 C=2 G=3
}
__global__ void node_ids_kernel(const uint64_t *__restrict__ table, uint64_t n_nodes,
                                const uint64_t *__restrict__ codes, uint64_t n, uint64_t *__restrict__ ids) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const NodeLookup L{nullptr, nullptr, table, n_nodes, 0};
    const int64_t id = find_code(L, codes[i]);
    ids[i] = id < 0 ? ~0ULL : (uint64_t)id;
}

int kmer_stats(pdg_graph *g, const uint64_t *h_begin, const uint64_t *h_len, uint64_t n, uint64_t *n_kmers,
               uint64_t *n_unique) {
    cudaStream_t s = g->stream;
    const int k = g->k;
    std::vector<uint64_t> begin, len;
    std::vector<uint32_t> type;
    for (uint64_t i = 0; i < n; ++i) {
        if (h_len[i] <= (uint64_t)k) continue;
        if (h_begin[i] + h_len[i] > g->n_res) {
            set_error("pdg_kmer_stats: contig beyond the residue store");
            return PDG_E_INVALID;
        }
        if (!begin.empty() && h_begin[i] < begin.back() + len.back()) {
            set_error("pdg_kmer_stats: contigs must be ascending and disjoint");
            return PDG_E_INVALID;
        }
        begin.push_back(h_begin[i]);
        len.push_back(h_len[i]);
        type.push_back(0);
    }
    const HostContigs hc(begin, len, type, k);
    *n_kmers = hc.n_kmers;
    *n_unique = 0;
    if (hc.n_kmers == 0) return PDG_OK;
    DevBuf db, de, dx, dt, dsum;
    const ContigTable ct = hc.upload(db, de, dx, dt, s);
    dsum.reserve(16, s);
    PDG_CUDA(cudaMemsetAsync(dsum.p, 0, 16, s));
    TraverseOut out;
    if (2 * k <= kBitmapMaxBits) {
        const uint64_t n_words = std::max<uint64_t>(1, (1ULL << (2 * k)) >> 5);
        DevBuf bm;
        bm.reserve(n_words * 4, s);
        PDG_CUDA(cudaMemsetAsync(bm.p, 0, n_words * 4, s));
        out.bitmap = bm.as<uint32_t>();
        run_traverse(g, MODE_MARK, ct, hc.pos0, hc.pos1, out);
        count_set_bits_kernel<<<grid_for(n_words, 256), 256, 0, s>>>(bm.as<uint32_t>(), n_words,
                                                                    dsum.as<unsigned long long>());
        PDG_LAUNCHED(g);
        PDG_CUDA(cudaMemcpyAsync(n_unique, dsum.p, 8, cudaMemcpyDeviceToHost, s));
        PDG_CUDA(cudaStreamSynchronize(s));
    } else {
        DevBuf ka, kb;
        ka.reserve(hc.n_kmers * 8, s);
        kb.reserve(hc.n_kmers * 8, s);
        out.codes = ka.as<uint64_t>();
        run_traverse(g, MODE_CODES, ct, hc.pos0, hc.pos1, out);
        uint64_t pr = 0;
        const uint64_t *sorted = radix_sort_keys(g, ka.as<uint64_t>(), kb.as<uint64_t>(), hc.n_kmers, 0, 2 * k, &pr);
        count_heads_kernel<<<grid_for(hc.n_kmers, 256), 256, 0, s>>>(sorted, hc.n_kmers, dsum.as<unsigned long long>());
        PDG_LAUNCHED(g);
        PDG_CUDA(cudaMemcpyAsync(n_unique, dsum.p, 8, cudaMemcpyDeviceToHost, s));
        PDG_CUDA(cudaStreamSynchronize(s));
    }
    return PDG_OK;
}

int emit_codes(pdg_graph *g, uint64_t begin, uint64_t len, uint64_t *d_out) {
    cudaStream_t s = g->stream;
    if (begin + len > g->n_res) {
        set_error("pdg_emit_codes: range beyond the residue store");
        return PDG_E_INVALID;
    }
    if (len < (uint64_t)g->k) return PDG_OK;
    // one pseudo contig; HostContigs wants len > k, so a single k-mer is padded by one and clipped
    std::vector<uint64_t> b{begin}, l{len};
    std::vector<uint32_t> t{0};
    HostContigs hc(b, l, t, g->k);
    DevBuf db, de, dx, dt;
    const ContigTable ct = hc.upload(db, de, dx, dt, s);
    TraverseOut out;
    out.codes = d_out;
    run_traverse(g, MODE_CODES, ct, begin, begin + len, out);
    PDG_CUDA(cudaStreamSynchronize(s));
    return PDG_OK;
}

int fetch_residues(pdg_graph *g, uint64_t begin, uint64_t len, uint8_t *h_ascii) {
    cudaStream_t s = g->stream;
    if (begin + len > g->n_res) {
        set_error("pdg_fetch_residues: range beyond the residue store");
        return PDG_E_INVALID;
    }
    if (len == 0) return PDG_OK;
    DevBuf d;
    d.reserve(len, s);
    decode_residues_kernel<<<grid_for(len, 256), 256, 0, s>>>(g->packed.as<unsigned long long>(), begin, len,
                                                             d.as<uint8_t>());
    PDG_LAUNCHED(g);
    PDG_CUDA(cudaMemcpyAsync(h_ascii, d.p, len, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    return PDG_OK;
}

int node_ids(pdg_graph *g, const uint64_t *h_codes, uint64_t n, uint64_t *h_ids) {
    cudaStream_t s = g->stream;
    if (g->u_lazy) {   // multi-rank: collective on the first call after pdg_finalize (before the n == 0 exit)
        const int rc = comm_materialize_nodes(g);
        if (rc != PDG_OK) return rc;
    }
    if (n == 0) return PDG_OK;
    DevBuf dc, di;
    dc.reserve(n * 8, s);
    di.reserve(n * 8, s);
    PDG_CUDA(cudaMemcpyAsync(dc.p, h_codes, n * 8, cudaMemcpyHostToDevice, s));
    node_ids_kernel<<<grid_for(n, 256), 256, 0, s>>>(g->u_codes, g->n_nodes, dc.as<uint64_t>(), n, di.as<uint64_t>());
    PDG_LAUNCHED(g);
    PDG_CUDA(cudaMemcpyAsync(h_ids, di.p, n * 8, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    return PDG_OK;
}

}  // namespace pdg
