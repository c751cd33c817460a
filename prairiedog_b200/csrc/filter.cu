// filter.cu -- Graph.filter() on the device.
//
// The only backend of the reference that implements filter() is NetworkXGraph (prairiedog/networkx_graph.py:53-75;
// LGGraph.filter is `pass`, lemon_graph.py:172-173):
//   :53-67  _filter_degree   max_degrees = max degree; bottom_degrees = int(0.9 * max_degrees);
//                            remove every node with degree <= bottom_degrees
//   :69-71  _filter_isolates remove every node that is left without a neighbour
// on an nx.Graph: UNDIRECTED, one edge per node pair whatever the labels, degree = number of distinct neighbours
// with a self-loop counting twice.  Here the labelled multigraph (genome part + explicit part) is projected onto that
// simple graph on the fly:
//   every labelled edge (src, tgt) -> the two keys (u << 32 | v), (v << 32 | u) over dense node ids -> radix sort +
//   run-length dedup = the adjacency of the simple graph -> degrees -> the two keep masks -> the labelled edges whose
//   endpoints both survive (the induced sub-multigraph), compacted for the host.
// tests/golden/filter.json holds what the reference's own methods, run unmodified on an nx.Graph, leave of the test
// graphs (tests/test_gpu_filter.py).
#include "gindex.cuh"
#include <algorithm>

namespace pdg {

int materialise_edges(pdg_graph *g, DevBuf &d_src, DevBuf &d_tgt, DevBuf &d_type, DevBuf &d_val);   // export.cu

// dense id of a node code in the sorted table of the whole graph
__device__ __forceinline__ uint32_t node_id_of(const uint64_t *__restrict__ codes, uint64_t n, uint64_t code) {
    uint64_t lo = 0, hi = n;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (codes[mid] < code) lo = mid + 1; else hi = mid;
    }
    return (uint32_t)lo;
}

// ids of both endpoints (kept for the later passes) and the two directed keys of the undirected pair
__global__ void flt_keys_kernel(const uint64_t *__restrict__ src, const uint64_t *__restrict__ tgt, uint64_t n_edges,
                                const uint64_t *__restrict__ codes, uint64_t n_nodes, uint32_t *__restrict__ uid,
                                uint32_t *__restrict__ vid, uint64_t *__restrict__ keys) {
    const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_edges) return;
    const uint32_t u = node_id_of(codes, n_nodes, src[e]), v = node_id_of(codes, n_nodes, tgt[e]);
    uid[e] = u;
    vid[e] = v;
    keys[2 * e] = ((uint64_t)u << 32) | v;
    keys[2 * e + 1] = ((uint64_t)v << 32) | u;
}

// one key per distinct (node, neighbour): degree as networkx counts it (a self-loop counts twice).
// keep == nullptr: over all nodes; else only pairs whose two nodes are both kept
__global__ void flt_degree_kernel(const uint64_t *__restrict__ keys, uint64_t n, const uint8_t *__restrict__ keep,
                                  uint32_t *__restrict__ deg) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t u = (uint32_t)(keys[i] >> 32), v = (uint32_t)keys[i];
    if (keep && !(keep[u] && keep[v])) return;
    atomicAdd(&deg[u], u == v ? 2u : 1u);
}

__global__ void flt_max_kernel(const uint32_t *__restrict__ deg, uint64_t n, uint32_t *__restrict__ mx) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t d = i < n ? deg[i] : 0u;
    for (int o = 16; o; o >>= 1) d = max(d, __shfl_xor_sync(0xffffffffu, d, o));
    if ((threadIdx.x & 31) == 0 && d) atomicMax(mx, d);
}

// keep[u] = deg[u] > bottom (:53-67) / keep[u] &= deg2[u] > 0 (:69-71); counts the survivors
__global__ void flt_keep_kernel(const uint32_t *__restrict__ deg, uint64_t n, uint32_t bottom, int second,
                                uint8_t *__restrict__ keep, unsigned long long *__restrict__ n_kept) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool k = false;
    if (i < n) {
        k = second ? (keep[i] && deg[i] > 0u) : (deg[i] > bottom);
        keep[i] = k ? 1 : 0;
    }
    const uint32_t m = __ballot_sync(0xffffffffu, k);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(n_kept, (unsigned long long)__popc(m));
}

__global__ void flt_edge_flags_kernel(const uint32_t *__restrict__ uid, const uint32_t *__restrict__ vid, uint64_t n,
                                      const uint8_t *__restrict__ keep, uint32_t *__restrict__ flags) {
    const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e < n) flags[e] = (keep[uid[e]] && keep[vid[e]]) ? 1u : 0u;
}

__global__ void flt_compact_kernel(const uint32_t *__restrict__ flags, const uint64_t *__restrict__ off, uint64_t n,
                                   const uint64_t *__restrict__ src, const uint64_t *__restrict__ tgt,
                                   const uint32_t *__restrict__ type, const int32_t *__restrict__ val,
                                   uint64_t *__restrict__ o_src, uint64_t *__restrict__ o_tgt,
                                   uint32_t *__restrict__ o_type, int32_t *__restrict__ o_val) {
    const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n || !flags[e]) return;
    const uint64_t o = off[e];
    o_src[o] = src[e];
    o_tgt[o] = tgt[e];
    o_type[o] = type[e];
    o_val[o] = val[e];
}

// steps 1-3: adjacency of the simple graph and its degrees; *max_degree for the host's int(0.9 * max_degrees)
int filter_degrees(pdg_graph *g, uint32_t *max_degree) {
    cudaStream_t s = g->stream;
    if (g->world > 1) {
        set_error("pdg_filter_degrees: not available in a multi-rank job (the edges of a node are spread over the ranks)");
        return PDG_E_INVALID;
    }
    g->flt_ready = false;
    const uint64_t ne = g->n_edges, nn = g->n_nodes;
    *max_degree = 0;
    g->flt_pairs = 0;
    g->flt_deg.reserve(std::max<size_t>(nn * 4, 16), s);
    PDG_CUDA(cudaMemsetAsync(g->flt_deg.p, 0, std::max<size_t>(nn * 4, 16), s));
    if (ne) {
        const int rc = materialise_edges(g, g->flt_src, g->flt_tgt, g->flt_type, g->flt_val);
        if (rc != PDG_OK) return rc;
        g->flt_uid.reserve(ne * 4, s);
        g->flt_vid.reserve(ne * 4, s);
        DevBuf keys, tmp;
        keys.reserve(2 * ne * 8, s);
        tmp.reserve(2 * ne * 8, s);
        flt_keys_kernel<<<grid_for(ne, 256), 256, 0, s>>>(g->flt_src.as<uint64_t>(), g->flt_tgt.as<uint64_t>(), ne, g->u_codes, nn,
                                                         g->flt_uid.as<uint32_t>(), g->flt_vid.as<uint32_t>(), keys.as<uint64_t>());
        PDG_LAUNCHED(g);
        g->flt_pairs = sort_unique_keys(g, keys.as<uint64_t>(), tmp.as<uint64_t>(), 2 * ne, 32 + std::max(bit_length(nn), 1),
                                        g->flt_adj);
        flt_degree_kernel<<<grid_for(g->flt_pairs, 256), 256, 0, s>>>(g->flt_adj.as<uint64_t>(), g->flt_pairs, nullptr,
                                                                     g->flt_deg.as<uint32_t>());
        PDG_LAUNCHED(g);
    }
    g->small.reserve(32, s);
    PDG_CUDA(cudaMemsetAsync(g->small.p, 0, 32, s));
    if (nn) {
        flt_max_kernel<<<grid_for(nn, 256), 256, 0, s>>>(g->flt_deg.as<uint32_t>(), nn, g->small.as<uint32_t>());
        PDG_LAUNCHED(g);
    }
    PDG_CUDA(cudaMemcpyAsync(max_degree, g->small.p, 4, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    g->flt_ready = true;
    g->flt_selected = false;
    return PDG_OK;
}

// steps 4-6: the two keep masks and the labelled edges between surviving nodes
int filter_select(pdg_graph *g, uint32_t bottom_degrees, uint64_t *n_kept_nodes, uint64_t *n_kept_edges) {
    cudaStream_t s = g->stream;
    if (!g->flt_ready) {
        set_error("pdg_filter_select before pdg_filter_degrees");
        return PDG_E_INVALID;
    }
    const uint64_t ne = g->n_edges, nn = g->n_nodes;
    *n_kept_nodes = *n_kept_edges = 0;
    g->flt_kept_edges = 0;
    g->flt_selected = true;
    if (nn == 0) return PDG_OK;
    DevBuf keep, deg2;
    keep.reserve(nn, s);
    deg2.reserve(nn * 4, s);
    g->small.reserve(32, s);
    PDG_CUDA(cudaMemsetAsync(g->small.p, 0, 32, s));
    PDG_CUDA(cudaMemsetAsync(deg2.p, 0, nn * 4, s));
    unsigned long long *cnt = g->small.as<unsigned long long>();
    flt_keep_kernel<<<grid_for(nn, 256), 256, 0, s>>>(g->flt_deg.as<uint32_t>(), nn, bottom_degrees, 0, keep.as<uint8_t>(), cnt);
    PDG_LAUNCHED(g);
    if (g->flt_pairs) {
        flt_degree_kernel<<<grid_for(g->flt_pairs, 256), 256, 0, s>>>(g->flt_adj.as<uint64_t>(), g->flt_pairs, keep.as<uint8_t>(),
                                                                     deg2.as<uint32_t>());
        PDG_LAUNCHED(g);
    }
    flt_keep_kernel<<<grid_for(nn, 256), 256, 0, s>>>(deg2.as<uint32_t>(), nn, 0u, 1, keep.as<uint8_t>(), cnt + 1);
    PDG_LAUNCHED(g);
    unsigned long long h[2] = {0, 0};
    uint64_t kept_e = 0;
    if (ne) {
        DevBuf flags, off;
        flags.reserve(ne * 4, s);
        off.reserve((ne + 1) * 8, s);
        flt_edge_flags_kernel<<<grid_for(ne, 256), 256, 0, s>>>(g->flt_uid.as<uint32_t>(), g->flt_vid.as<uint32_t>(), ne,
                                                               keep.as<uint8_t>(), flags.as<uint32_t>());
        PDG_LAUNCHED(g);
        exclusive_scan_u32_u64(flags.as<uint32_t>(), off.as<uint64_t>(), ne, g);
        PDG_CUDA(cudaMemcpyAsync(&kept_e, off.as<uint64_t>() + ne, 8, cudaMemcpyDeviceToHost, s));
        PDG_CUDA(cudaStreamSynchronize(s));
        if (kept_e) {
            g->flt_osrc.reserve(kept_e * 8, s);
            g->flt_otgt.reserve(kept_e * 8, s);
            g->flt_otype.reserve(kept_e * 4, s);
            g->flt_oval.reserve(kept_e * 4, s);
            flt_compact_kernel<<<grid_for(ne, 256), 256, 0, s>>>(
                flags.as<uint32_t>(), off.as<uint64_t>(), ne, g->flt_src.as<uint64_t>(), g->flt_tgt.as<uint64_t>(),
                g->flt_type.as<uint32_t>(), g->flt_val.as<int32_t>(), g->flt_osrc.as<uint64_t>(), g->flt_otgt.as<uint64_t>(),
                g->flt_otype.as<uint32_t>(), g->flt_oval.as<int32_t>());
            PDG_LAUNCHED(g);
        }
        PDG_CUDA(cudaStreamSynchronize(s));    // flags / off die with this scope
    }
    PDG_CUDA(cudaMemcpyAsync(h, cnt, 16, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    g->flt_kept_edges = kept_e;
    *n_kept_nodes = h[1];
    *n_kept_edges = kept_e;
    return PDG_OK;
}

int filter_fetch(pdg_graph *g, uint64_t *h_src, uint64_t *h_tgt, uint32_t *h_type, int32_t *h_val) {
    cudaStream_t s = g->stream;
    if (!g->flt_ready || !g->flt_selected) {
        set_error("pdg_filter_fetch before pdg_filter_select");
        return PDG_E_INVALID;
    }
    const uint64_t n = g->flt_kept_edges;
    if (n) {
        if (h_src) PDG_CUDA(cudaMemcpyAsync(h_src, g->flt_osrc.p, n * 8, cudaMemcpyDeviceToHost, s));
        if (h_tgt) PDG_CUDA(cudaMemcpyAsync(h_tgt, g->flt_otgt.p, n * 8, cudaMemcpyDeviceToHost, s));
        if (h_type) PDG_CUDA(cudaMemcpyAsync(h_type, g->flt_otype.p, n * 4, cudaMemcpyDeviceToHost, s));
        if (h_val) PDG_CUDA(cudaMemcpyAsync(h_val, g->flt_oval.p, n * 4, cudaMemcpyDeviceToHost, s));
    }
    PDG_CUDA(cudaStreamSynchronize(s));
    // the working set is only needed between the three calls
    g->flt_src.release(); g->flt_tgt.release(); g->flt_type.release(); g->flt_val.release();
    g->flt_uid.release(); g->flt_vid.release(); g->flt_adj.release(); g->flt_deg.release();
    g->flt_osrc.release(); g->flt_otgt.release(); g->flt_otype.release(); g->flt_oval.release();
    g->flt_ready = g->flt_selected = false;
    return PDG_OK;
}

}  // namespace pdg
