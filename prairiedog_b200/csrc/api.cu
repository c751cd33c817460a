// api.cu -- the extern "C" surface declared in include/pdgpu.h.
#include "common.cuh"
#include <stdarg.h>
#include <string.h>
#include <new>
#include <algorithm>

namespace pdg {
static thread_local char g_err[1024] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace pdg

using namespace pdg;

#define PDG_API_BEGIN                                   \
    if (!g) {                                           \
        set_error("null graph handle");                 \
        return PDG_E_INVALID;                           \
    }                                                   \
    try {                                               \
        if (cudaSetDevice(g->device) != cudaSuccess) {  \
            set_error("cudaSetDevice(%d) failed", g->device); \
            return PDG_E_CUDA;                          \
        }
#define PDG_API_END                                     \
    }                                                   \
    catch (const pdg::CudaError &e) { return e.code; }  \
    catch (const std::bad_alloc &) {                    \
        set_error("host allocation failed");            \
        return PDG_E_NOMEM;                             \
    }

extern "C" {

const char *pdg_last_error(void) { return g_err; }
int pdg_abi_version(void) { return PDG_ABI_VERSION; }
int pdg_struct_sizes(uint64_t *hit, uint64_t *ingest_info, uint64_t *build_stats) {
    if (hit) *hit = sizeof(pdg_hit);
    if (ingest_info) *ingest_info = sizeof(pdg_ingest_info);
    if (build_stats) *build_stats = sizeof(pdg_build_stats);
    return PDG_OK;
}

pdg_graph *pdg_create(int device, int k) {
    if (k < 1 || k > 31) {
        set_error("pdg_create: k=%d outside [1,31]", k);
        return nullptr;
    }
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) {
        set_error("pdg_create: CUDA device %d not available (%d devices) -- there is no CPU path", device, n);
        return nullptr;
    }
    if (cudaSetDevice(device) != cudaSuccess) {
        set_error("pdg_create: cudaSetDevice(%d) failed", device);
        return nullptr;
    }
    pdg_graph *g = new (std::nothrow) pdg_graph();
    if (!g) return nullptr;
    g->device = device;
    g->k = k;
    return g;
}

void pdg_destroy(pdg_graph *g) {
    if (!g) return;
    cudaSetDevice(g->device);
    cudaStreamSynchronize(g->stream);
    comm_destroy(g);
    if (g->ev0) cudaEventDestroy(g->ev0);
    if (g->ev1) cudaEventDestroy(g->ev1);
    if (g->fb_ev_ready) cudaEventDestroy(g->fb_ev_ready);
    for (cudaEvent_t e : g->fb_events) cudaEventDestroy(e);
    for (cudaEvent_t e : g->k_events) cudaEventDestroy(e);
    for (cudaEvent_t e : g->i_events) cudaEventDestroy(e);
    if (g->copy_stream) {
        cudaStreamSynchronize(g->copy_stream);
        cudaStreamDestroy(g->copy_stream);
    }
    if (g->aux_stream) {
        cudaStreamSynchronize(g->aux_stream);
        cudaStreamDestroy(g->aux_stream);
    }
    delete g;
}

int pdg_set_stream(pdg_graph *g, void *cuda_stream) {
    PDG_API_BEGIN
    g->stream = reinterpret_cast<cudaStream_t>(cuda_stream);
    return PDG_OK;
    PDG_API_END
}

int pdg_clear(pdg_graph *g) {
    PDG_API_BEGIN
    g->pre_valid = false;   // scratch buffers of a pipelined ingest may be reused from here on
    g->m3_groups.clear();
    g->m3_launched = false;
    PDG_CUDA(cudaStreamSynchronize(g->stream));
    g->n_res = 0;
    if (g->packed.p) PDG_CUDA(cudaMemsetAsync(g->packed.p, 0, g->packed.cap, g->stream));
    g->c_begin.clear(); g->c_len.clear(); g->c_type.clear();
    g->g_begin.clear(); g->g_len.clear(); g->g_type.clear();
    g->n_kmers = 0;
    g->x_src.clear(); g->x_tgt.clear(); g->x_nodes.clear(); g->x_type.clear(); g->x_val.clear();
    g->finalized = false;
    g->n_nodes = g->n_edges = 0;
    g->g_kmers = g->g_edges = g->g_nodes = 0;
    g->x_n_nodes = g->x_n_edges = 0;
    g->q_n = g->q_total = 0;
    g->flt_ready = g->flt_selected = false;
    g->launches = 0;
    return PDG_OK;
    PDG_API_END
}

int pdg_ingest_fasta(pdg_graph *g, const uint8_t *d_bytes, uint64_t n_bytes, uint64_t *h_header_table,
                     uint64_t cap_headers, pdg_ingest_info *info) {
    PDG_API_BEGIN
    g->pre_valid = false;   // scratch buffers of a pipelined ingest may be reused from here on
    if (!info || (n_bytes && !d_bytes)) {
        set_error("pdg_ingest_fasta: null argument");
        return PDG_E_INVALID;
    }
    return ingest_fasta(g, d_bytes, n_bytes, h_header_table, cap_headers, info);
    PDG_API_END
}

int pdg_ingest_fasta_batch(pdg_graph *g, const uint8_t *const *ptrs, const uint64_t *n_bytes, uint32_t n_files,
                           int flags, uint64_t *h_header_table, uint64_t cap_headers, pdg_ingest_info *infos) {
    PDG_API_BEGIN
    if (n_files && (!ptrs || !n_bytes || !infos)) {
        set_error("pdg_ingest_fasta_batch: null argument");
        return PDG_E_INVALID;
    }
    return ingest_fasta_batch(g, ptrs, n_bytes, n_files, flags, h_header_table, cap_headers, infos);
    PDG_API_END
}

int pdg_add_contigs(pdg_graph *g, const uint64_t *h_begin, const uint64_t *h_len, const uint32_t *h_type_id,
                    uint64_t n_contigs) {
    PDG_API_BEGIN
    const uint64_t k = (uint64_t)g->k;
    for (uint64_t i = 0; i < n_contigs; ++i) {
        if (h_begin[i] + h_len[i] > g->n_res) {
            set_error("pdg_add_contigs: contig %llu [%llu,+%llu) beyond the %llu residues in the store",
                      (unsigned long long)i, (unsigned long long)h_begin[i], (unsigned long long)h_len[i],
                      (unsigned long long)g->n_res);
            return PDG_E_INVALID;
        }
        if (h_len[i] <= k) continue;   // kmers.py:110-117: a contig with len <= k yields nothing
        if (!g->c_begin.empty() && h_begin[i] < g->c_begin.back() + g->c_len.back()) {
            set_error("pdg_add_contigs: contigs must be ascending and disjoint");
            return PDG_E_INVALID;
        }
        g->c_begin.push_back(h_begin[i]);
        g->c_len.push_back(h_len[i]);
        g->c_type.push_back(h_type_id[i]);
        g->n_kmers += h_len[i] - k + 1;
    }
    // g->finalized stays as it is: declaring contigs only appends to host vectors, the index of the last
    // pdg_finalize is untouched and keeps answering queries -- LemonGraph's read transactions see the last COMMITTED
    // state while a write transaction is open (prairiedog/lemon_graph.py:178-180 against :54-64)
    return PDG_OK;
    PDG_API_END
}

int pdg_kmer_stats(pdg_graph *g, const uint64_t *h_begin, const uint64_t *h_len, uint64_t n_contigs,
                   uint64_t *n_kmers, uint64_t *n_unique) {
    PDG_API_BEGIN
    g->pre_valid = false;   // scratch buffers of a pipelined ingest may be reused from here on
    return kmer_stats(g, h_begin, h_len, n_contigs, n_kmers, n_unique);
    PDG_API_END
}

int pdg_emit_codes(pdg_graph *g, uint64_t begin, uint64_t len, uint64_t *d_codes_out) {
    PDG_API_BEGIN
    g->pre_valid = false;   // scratch buffers of a pipelined ingest may be reused from here on
    return emit_codes(g, begin, len, d_codes_out);
    PDG_API_END
}

int pdg_fetch_residues(pdg_graph *g, uint64_t begin, uint64_t len, uint8_t *h_ascii) {
    PDG_API_BEGIN
    return fetch_residues(g, begin, len, h_ascii);
    PDG_API_END
}

int pdg_snapshot_sizes(pdg_graph *g, uint64_t *out4) {
    PDG_API_BEGIN
    out4[0] = g->n_res;
    out4[1] = g->c_begin.size();
    out4[2] = g->x_src.size();
    out4[3] = g->x_nodes.size();
    return PDG_OK;
    PDG_API_END
}

int pdg_snapshot_export(pdg_graph *g, uint64_t *h_words, uint64_t cap_words, uint64_t *h_begin, uint64_t *h_len,
                        uint32_t *h_type_id, uint64_t *h_src_code, uint64_t *h_tgt_code, uint32_t *h_edge_type_id,
                        int32_t *h_value, uint64_t *h_node_code) {
    PDG_API_BEGIN
    const uint64_t words = (g->n_res + 31) / 32;
    if ((words && !h_words) || (!g->c_begin.empty() && (!h_begin || !h_len || !h_type_id)) ||
        (!g->x_src.empty() && (!h_src_code || !h_tgt_code || !h_edge_type_id || !h_value)) ||
        (!g->x_nodes.empty() && !h_node_code)) {
        set_error("pdg_snapshot_export: null output for a non-empty part (sizes from pdg_snapshot_sizes)");
        return PDG_E_INVALID;
    }
    if (words > cap_words) {
        set_error("pdg_snapshot_export: %llu words, capacity %llu", (unsigned long long)words,
                  (unsigned long long)cap_words);
        return PDG_E_CAPACITY;
    }
    if (words) {
        PDG_CUDA(cudaMemcpyAsync(h_words, g->packed.p, words * 8, cudaMemcpyDeviceToHost, g->stream));
        PDG_CUDA(cudaStreamSynchronize(g->stream));
    }
    std::copy(g->c_begin.begin(), g->c_begin.end(), h_begin);
    std::copy(g->c_len.begin(), g->c_len.end(), h_len);
    std::copy(g->c_type.begin(), g->c_type.end(), h_type_id);
    std::copy(g->x_src.begin(), g->x_src.end(), h_src_code);
    std::copy(g->x_tgt.begin(), g->x_tgt.end(), h_tgt_code);
    std::copy(g->x_type.begin(), g->x_type.end(), h_edge_type_id);
    std::copy(g->x_val.begin(), g->x_val.end(), h_value);
    std::copy(g->x_nodes.begin(), g->x_nodes.end(), h_node_code);
    return PDG_OK;
    PDG_API_END
}

int pdg_store_import(pdg_graph *g, const uint64_t *h_words, uint64_t n_residues) {
    PDG_API_BEGIN
    if (g->n_res || !g->c_begin.empty()) {
        set_error("pdg_store_import: the graph already holds %llu residues (pdg_clear first)",
                  (unsigned long long)g->n_res);
        return PDG_E_INVALID;
    }
    if (n_residues && !h_words) {
        set_error("pdg_store_import: null store");
        return PDG_E_INVALID;
    }
    if (n_residues >= (1ULL << 32)) {
        set_error("pdg_store_import: %llu residues exceed the 32-bit position column", (unsigned long long)n_residues);
        return PDG_E_INVALID;
    }
    g->pre_valid = false;
    const uint64_t words = (n_residues + 31) / 32;
    g->packed.reserve((words + 2) * 8, g->stream, /*keep=*/false, /*zero=*/true);
    PDG_CUDA(cudaMemsetAsync(g->packed.p, 0, g->packed.cap, g->stream));
    if (words) {
        PDG_CUDA(cudaMemcpyAsync(g->packed.p, h_words, words * 8, cudaMemcpyHostToDevice, g->stream));
        // the tail of the last word must be zero: later appends OR into it
        const uint32_t used = (uint32_t)(n_residues & 31);
        if (used) {
            const uint64_t last = h_words[words - 1] & ~((1ULL << (64 - 2 * used)) - 1ULL);
            PDG_CUDA(cudaStreamSynchronize(g->stream));
            PDG_CUDA(cudaMemcpyAsync(g->packed.as<uint64_t>() + words - 1, &last, 8, cudaMemcpyHostToDevice, g->stream));
        }
        PDG_CUDA(cudaStreamSynchronize(g->stream));
    }
    g->n_res = n_residues;
    g->finalized = false;
    return PDG_OK;
    PDG_API_END
}

int pdg_add_edges_host(pdg_graph *g, const uint64_t *h_src_code, const uint64_t *h_tgt_code,
                       const uint32_t *h_type_id, const int32_t *h_value, uint64_t n) {
    PDG_API_BEGIN
    g->x_src.insert(g->x_src.end(), h_src_code, h_src_code + n);
    g->x_tgt.insert(g->x_tgt.end(), h_tgt_code, h_tgt_code + n);
    g->x_type.insert(g->x_type.end(), h_type_id, h_type_id + n);
    g->x_val.insert(g->x_val.end(), h_value, h_value + n);
    return PDG_OK;      // committed index stays valid, see pdg_add_contigs
    PDG_API_END
}

int pdg_add_nodes_host(pdg_graph *g, const uint64_t *h_code, uint64_t n) {
    PDG_API_BEGIN
    g->x_nodes.insert(g->x_nodes.end(), h_code, h_code + n);
    return PDG_OK;      // committed index stays valid, see pdg_add_contigs
    PDG_API_END
}

int pdg_finalize(pdg_graph *g, int flags) {
    PDG_API_BEGIN
    return finalize(g, flags);
    PDG_API_END
}

int pdg_counts(pdg_graph *g, uint64_t *n_nodes, uint64_t *n_edges, uint64_t *n_kmers) {
    PDG_API_BEGIN
    if (n_nodes) *n_nodes = g->n_nodes;
    if (n_edges) *n_edges = g->n_edges;
    // after a pdg_finalize: the k-mers of the committed graph (contigs declared since are not in it yet)
    if (n_kmers) *n_kmers = g->finalized ? g->stats.n_kmers : g->n_kmers;
    return PDG_OK;
    PDG_API_END
}

int pdg_get_build_stats(pdg_graph *g, pdg_build_stats *out) {
    PDG_API_BEGIN
    *out = g->stats;
    out->kernel_launches = g->launches;
    out->ms_ingest_a = g->ms_ingest[0];
    out->ms_ingest_b = g->ms_ingest[1];
    out->ms_ingest_c = g->ms_ingest[2];
    out->ingest_bytes = g->ingest_bytes;
    out->ingest_launches = g->ingest_launches;
    return PDG_OK;
    PDG_API_END
}

static int need_final(pdg_graph *g, const char *who) {
    if (!g->finalized) {
        set_error("%s before pdg_finalize", who);
        return PDG_E_INVALID;
    }
    return PDG_OK;
}

int pdg_export_nodes(pdg_graph *g, uint64_t *h_codes) {
    PDG_API_BEGIN
    if (need_final(g, "pdg_export_nodes")) return PDG_E_INVALID;
    if (g->u_lazy) {   // multi-rank: collective on the first call after pdg_finalize
        const int rc = comm_materialize_nodes(g);
        if (rc != PDG_OK) return rc;
    }
    if (h_codes && g->n_nodes)
        PDG_CUDA(cudaMemcpyAsync(h_codes, g->u_codes, g->n_nodes * 8, cudaMemcpyDeviceToHost, g->stream));
    PDG_CUDA(cudaStreamSynchronize(g->stream));
    return PDG_OK;
    PDG_API_END
}

int pdg_export_edges(pdg_graph *g, uint64_t *h_src_code, uint64_t *h_tgt_code, uint32_t *h_type_id,
                     int32_t *h_value) {
    PDG_API_BEGIN
    if (need_final(g, "pdg_export_edges")) return PDG_E_INVALID;
    return export_edges(g, h_src_code, h_tgt_code, h_type_id, h_value);
    PDG_API_END
}

int pdg_node_ids(pdg_graph *g, const uint64_t *h_codes, uint64_t n, uint64_t *h_ids) {
    PDG_API_BEGIN
    if (need_final(g, "pdg_node_ids")) return PDG_E_INVALID;
    return node_ids(g, h_codes, n, h_ids);
    PDG_API_END
}

int pdg_index_size(pdg_graph *g, uint64_t *n_rows, uint64_t *n_positions) {
    PDG_API_BEGIN
    if (need_final(g, "pdg_index_size")) return PDG_E_INVALID;
    if (n_rows) *n_rows = g->g_nodes;
    if (n_positions) *n_positions = g->g_kmers;
    return PDG_OK;
    PDG_API_END
}

int pdg_export_index(pdg_graph *g, uint64_t *h_row_codes, uint32_t *h_row_ptr, uint32_t *h_positions) {
    PDG_API_BEGIN
    if (need_final(g, "pdg_export_index")) return PDG_E_INVALID;
    cudaStream_t s = g->stream;
    if (h_row_codes && g->g_nodes)
        PDG_CUDA(cudaMemcpyAsync(h_row_codes, g->g_codes.p, g->g_nodes * 8, cudaMemcpyDeviceToHost, s));
    if (h_row_ptr)
        PDG_CUDA(cudaMemcpyAsync(h_row_ptr, g->g_rowptr.p, (g->g_nodes + 1) * 4, cudaMemcpyDeviceToHost, s));
    if (h_positions && g->g_kmers)
        PDG_CUDA(cudaMemcpyAsync(h_positions, g->g_pos.p, g->g_kmers * 4, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    return PDG_OK;
    PDG_API_END
}

int pdg_checksum(pdg_graph *g, uint64_t *nodes_sum, uint64_t *edges_sum) {
    PDG_API_BEGIN
    if (need_final(g, "pdg_checksum")) return PDG_E_INVALID;
    return checksum(g, nodes_sum, edges_sum);
    PDG_API_END
}

int pdg_query_connected(pdg_graph *g, const uint64_t *h_a, const uint64_t *h_b, uint64_t n_q,
                        uint64_t *total_starts) {
    PDG_API_BEGIN
    return query_run(g, h_a, h_b, n_q, -1, true, total_starts);
    PDG_API_END
}

int pdg_query_paths(pdg_graph *g, const uint64_t *h_a, const uint64_t *h_b, uint64_t n_q, int64_t max_len,
                    uint64_t *total_hits) {
    PDG_API_BEGIN
    return query_run(g, h_a, h_b, n_q, max_len, false, total_hits);
    PDG_API_END
}

int pdg_query_fetch(pdg_graph *g, uint64_t *h_offsets, pdg_hit *h_hits) {
    PDG_API_BEGIN
    cudaStream_t s = g->stream;
    // cudaMemcpyDefault: the destinations may be host OR device memory (unified addressing tells them apart) -- a
    // multi-rank caller keeps the hits on the device for the exchange between the ranks
    if (h_offsets) PDG_CUDA(cudaMemcpyAsync(h_offsets, g->q_off.p, (g->q_n + 1) * 8, cudaMemcpyDefault, s));
    if (h_hits && g->q_total)
        PDG_CUDA(cudaMemcpyAsync(h_hits, g->q_hits.p, g->q_total * sizeof(pdg_hit), cudaMemcpyDefault, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    return PDG_OK;
    PDG_API_END
}

int pdg_query_walk(pdg_graph *g, uint64_t src_code, const pdg_hit *hit, uint64_t *h_node_codes, int32_t *status) {
    PDG_API_BEGIN
    return query_walk(g, src_code, hit, h_node_codes, status);
    PDG_API_END
}

int pdg_filter_degrees(pdg_graph *g, uint32_t *max_degree) {
    PDG_API_BEGIN
    if (need_final(g, "pdg_filter_degrees")) return PDG_E_INVALID;
    if (!max_degree) {
        set_error("pdg_filter_degrees: null argument");
        return PDG_E_INVALID;
    }
    return filter_degrees(g, max_degree);
    PDG_API_END
}

int pdg_filter_select(pdg_graph *g, uint32_t bottom_degrees, uint64_t *n_kept_nodes, uint64_t *n_kept_edges) {
    PDG_API_BEGIN
    if (need_final(g, "pdg_filter_select")) return PDG_E_INVALID;
    if (!n_kept_nodes || !n_kept_edges) {
        set_error("pdg_filter_select: null argument");
        return PDG_E_INVALID;
    }
    return filter_select(g, bottom_degrees, n_kept_nodes, n_kept_edges);
    PDG_API_END
}

int pdg_filter_fetch(pdg_graph *g, uint64_t *h_src_code, uint64_t *h_tgt_code, uint32_t *h_type_id, int32_t *h_value) {
    PDG_API_BEGIN
    return filter_fetch(g, h_src_code, h_tgt_code, h_type_id, h_value);
    PDG_API_END
}

int pdg_query_strings_size(pdg_graph *g, uint64_t *total_bytes) {
    PDG_API_BEGIN
    if (!total_bytes) {
        set_error("pdg_query_strings_size: null argument");
        return PDG_E_INVALID;
    }
    return query_strings_size(g, total_bytes);
    PDG_API_END
}

int pdg_query_strings_fetch(pdg_graph *g, uint64_t *h_offsets, uint8_t *h_bytes) {
    PDG_API_BEGIN
    return query_strings_fetch(g, h_offsets, h_bytes);
    PDG_API_END
}

}  // extern "C"
