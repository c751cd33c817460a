// build.cu -- pdg_finalize: splits the declared contigs into the genome part (occurrence index,
// gbuild.cu) and the explicit part (xbuild.cu), then forms the node table of the whole graph.
//
//   prairiedog/subgraph_ref.py:39-115     SubgraphRef.update_graph (the loop this replaces)
//   prairiedog/lemon_graph.py:154-157     LGGraph.save (publish)
//   lemongraph/lib/lemongraph.c:674-685   edges are unique on (type, value, src, tgt): a type id
//                                         that labels more than one contig, or a contig and an
//                                         explicit edge, could repeat a (type, value) pair, so
//                                         those contigs are expanded into the explicit part
//                                         where a full-key sort removes duplicates
#include "gindex.cuh"
#include <algorithm>
#include <chrono>
#include <unordered_map>

namespace pdg {

struct Timer {
    cudaStream_t s;
    std::chrono::steady_clock::time_point t0;
    explicit Timer(cudaStream_t st) : s(st) { cudaStreamSynchronize(s); t0 = std::chrono::steady_clock::now(); }
    float lap() {
        cudaStreamSynchronize(s);
        auto t1 = std::chrono::steady_clock::now();
        float ms = std::chrono::duration<float, std::milli>(t1 - t0).count();
        t0 = t1;
        return ms;
    }
};

int finalize(pdg_graph *g, int flags) {
    cudaStream_t s = g->stream;
    Timer total(s), lap(s);
    g->stats = pdg_build_stats{};
    g->finalized = false;
    g->u_lazy = false;
    const int k = g->k;
    if (g->n_res >= (1ULL << 32)) {
        set_error("pdg_finalize: %llu residues exceed the 32-bit position column", (unsigned long long)g->n_res);
        return PDG_E_INVALID;
    }
    // ---- which contigs stay in G ----
    std::unordered_map<uint32_t, uint32_t> uses;   // type id -> labelled contigs (+2 if explicit)
    bool all_fresh = !(flags & PDG_F_FORCE_GENERAL) && g->x_type.empty();
    if (all_fresh)   // common case: strictly increasing type ids need no table
        for (size_t i = 1; i < g->c_type.size() && all_fresh; ++i) all_fresh = g->c_type[i] > g->c_type[i - 1];
    std::vector<uint64_t> db, dl;
    std::vector<uint32_t> dt;
    g->g_begin.clear(); g->g_len.clear(); g->g_type.clear();
    if (all_fresh) {
        g->g_begin = g->c_begin; g->g_len = g->c_len; g->g_type = g->c_type;
    } else {
        for (uint32_t t : g->c_type) uses[t] += 1;
        for (uint32_t t : g->x_type) uses[t] |= 0x40000000u;
        for (size_t i = 0; i < g->c_begin.size(); ++i) {
            const bool demote = (flags & PDG_F_FORCE_GENERAL) || uses[g->c_type[i]] != 1;
            (demote ? db : g->g_begin).push_back(g->c_begin[i]);
            (demote ? dl : g->g_len).push_back(g->c_len[i]);
            (demote ? dt : g->g_type).push_back(g->c_type[i]);
        }
    }
    const HostContigs hg(g->g_begin, g->g_len, g->g_type, k), hd(db, dl, dt, k);
    const ContigTable ct = hg.upload(g->d_gbegin, g->d_gend, g->d_gkbase, g->d_gtype, s);
    g->stats.n_demoted = hd.size();
    g->stats.ms_upload = lap.lap();

    // ---- G: occurrence index ----
    build_g_index(g, ct, hg, flags);
    g->stats.node_path = g->have_bitmap ? 0 : 1;
    g->stats.ms_sort = lap.lap();

    // ---- X: explicit part ----
    build_x_part(g, hd);
    g->stats.n_edge_records = g->x_src.size() + (hd.n_kmers - hd.size());
    g->stats.ms_explicit = lap.lap();

    // ---- node table of the whole graph ----
    uint64_t n_local = 0;
    if (g->x_n_nodes == 0) {
        g->u_codes = g->g_codes.as<uint64_t>();
        n_local = g->g_nodes;
    } else if (g->g_nodes == 0) {
        g->u_codes = g->x_codes.as<uint64_t>();
        n_local = g->x_n_nodes;
    } else {
        const uint64_t n = g->g_nodes + g->x_n_nodes;
        DevBuf a, b;
        a.reserve(n * 8, s);
        b.reserve(n * 8, s);
        PDG_CUDA(cudaMemcpyAsync(a.p, g->g_codes.p, g->g_nodes * 8, cudaMemcpyDeviceToDevice, s));
        PDG_CUDA(cudaMemcpyAsync(a.as<uint64_t>() + g->g_nodes, g->x_codes.p, g->x_n_nodes * 8,
                                 cudaMemcpyDeviceToDevice, s));
        n_local = sort_unique_keys(g, a.as<uint64_t>(), b.as<uint64_t>(), n, 64, g->u_codes_buf);
        g->u_codes = g->u_codes_buf.as<uint64_t>();
        PDG_CUDA(cudaStreamSynchronize(s));
    }
    g->n_nodes = n_local;
    if (g->world > 1) {
        // K5: range-partitioned exchange of the node keys; every rank learns the global node count, the
        // global table is gathered on demand (comm_materialize_nodes)
        uint64_t n_glob = 0;
        // opaque node keys of the explicit part carry bit 63; clean k-mer codes need 2k bits
        const int rc = comm_global_nodes(g, g->u_codes, n_local, g->x_n_nodes ? 64 : 2 * k, &n_glob);
        if (rc != PDG_OK) return rc;
        g->u_codes = g->u_codes_buf.as<uint64_t>();
        g->n_nodes = n_glob;
    }
    g->n_edges = g->g_edges + g->x_n_edges;
    g->stats.ms_nodes = lap.lap();

    pdg_build_stats &st = g->stats;
    st.ms_total = total.lap();
    st.n_residues = g->n_res;
    st.n_contigs = g->c_begin.size();
    st.n_kmers = g->n_kmers;
    st.n_nodes = g->n_nodes;
    st.n_edges = g->n_edges;
    st.kernel_launches = g->launches;
    g->finalized = true;
    return PDG_OK;
}

}  // namespace pdg
