// query.cu -- K4: batched "all paths between two k-mers" over the CSR/CSC.
//
// Restates LGGraph.connected / path (prairiedog/lemon_graph.py:178-334) and Graph.matching_edges
// (prairiedog/graph.py:82-113).  Out-edges of a (CSR row) and in-edges of b (CSC row) are both
// sorted by (type id, value); a start edge (T,i) matches every in-edge (T,j) with i <= j
// (and j - i <= max_len for path()), which is one contiguous range found by binary search.
// One warp per query; a count launch, a prefix sum, then a fill launch.
#include "common.cuh"

namespace pdg {

struct GraphView {
    const uint64_t *codes;
    uint64_t n_nodes;
    const uint64_t *row_ptr;
    const uint32_t *csr_tgt, *csr_type, *csr_val;
    const uint64_t *col_ptr;
    const uint32_t *csc_src, *csc_type, *csc_val;
};

static GraphView view_of(const pdg_graph *g) {
    GraphView v;
    v.codes = g->node_codes.as<uint64_t>();
    v.n_nodes = g->n_nodes;
    v.row_ptr = g->row_ptr.as<uint64_t>();
    v.csr_tgt = g->csr_tgt.as<uint32_t>();
    v.csr_type = g->csr_type.as<uint32_t>();
    v.csr_val = g->csr_val.as<uint32_t>();
    v.col_ptr = g->col_ptr.as<uint64_t>();
    v.csc_src = g->csc_src.as<uint32_t>();
    v.csc_type = g->csc_type.as<uint32_t>();
    v.csc_val = g->csc_val.as<uint32_t>();
    return v;
}

__device__ __forceinline__ int64_t find_node(const GraphView &G, uint64_t code) {
    uint64_t lo = 0, hi = G.n_nodes;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (G.codes[mid] < code) lo = mid + 1; else hi = mid;
    }
    return (lo < G.n_nodes && G.codes[lo] == code) ? (int64_t)lo : -1;
}

__device__ __forceinline__ uint64_t tv_key(const uint32_t *type, const uint32_t *val, uint64_t e) {
    return ((uint64_t)type[e] << 32) | (uint64_t)val[e];
}
// first index in [lo,hi) whose (type,val) key is >= key
__device__ __forceinline__ uint64_t lower_bound_tv(const uint32_t *type, const uint32_t *val, uint64_t lo, uint64_t hi,
                                                   uint64_t key) {
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (tv_key(type, val, mid) < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}
// first index in [lo,hi) whose key is > key
__device__ __forceinline__ uint64_t upper_bound_tv(const uint32_t *type, const uint32_t *val, uint64_t lo, uint64_t hi,
                                                   uint64_t key) {
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (tv_key(type, val, mid) <= key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// max_len < 0: no limit.  connected: count / emit start edges instead of (start,end) pairs.
template <bool FILL>
__global__ void __launch_bounds__(256) query_kernel(GraphView G, const uint64_t *__restrict__ qa,
                                                    const uint64_t *__restrict__ qb, uint64_t n_q, int64_t max_len,
                                                    int connected, uint32_t *__restrict__ counts,
                                                    const uint64_t *__restrict__ offsets, pdg_hit *__restrict__ hits) {
    const uint64_t q = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (q >= n_q) return;
    const int64_t ia = find_node(G, qa[q]);
    const int64_t ib = find_node(G, qb[q]);
    uint64_t total = 0;
    if (ia >= 0 && ib >= 0) {
        const uint64_t o0 = G.row_ptr[ia], o1 = G.row_ptr[ia + 1];
        const uint64_t i0 = G.col_ptr[ib], i1 = G.col_ptr[ib + 1];
        uint64_t out_base = FILL ? offsets[q] : 0;
        if (i1 > i0) {
            for (uint64_t base = o0; base < o1; base += 32) {
                const uint64_t e = base + lane;
                uint64_t lo = 0, cnt = 0;
                uint32_t ty = 0, vb = 0;
                if (e < o1) {
                    ty = G.csr_type[e];
                    vb = G.csr_val[e];
                    const uint64_t key_lo = ((uint64_t)ty << 32) | vb;
                    uint64_t vmax = 0xFFFFFFFFull;
                    if (!connected && max_len >= 0) vmax = min((uint64_t)vb + (uint64_t)max_len, (uint64_t)0xFFFFFFFFull);
                    const uint64_t key_hi = ((uint64_t)ty << 32) | vmax;
                    lo = lower_bound_tv(G.csc_type, G.csc_val, i0, i1, key_lo);
                    const uint64_t hi = upper_bound_tv(G.csc_type, G.csc_val, lo, i1, key_hi);
                    cnt = hi - lo;
                    if (connected) cnt = cnt ? 1 : 0;
                }
                // warp exclusive scan of cnt
                uint64_t inc = cnt;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint64_t t = __shfl_up_sync(0xffffffffu, inc, o);
                    if (lane >= o) inc += t;
                }
                const uint64_t wtot = __shfl_sync(0xffffffffu, inc, 31);
                if (FILL && cnt) {
                    pdg_hit *dst = hits + out_base + (inc - cnt);
                    for (uint64_t t = 0; t < cnt; ++t) {
                        pdg_hit h;
                        h.type_id = ty;
                        h.i = (int32_t)(vb ^ 0x80000000u);
                        h.j = connected ? h.i : (int32_t)(G.csc_val[lo + t] ^ 0x80000000u);
                        h.src_tgt_id = G.csr_tgt[e];
                        dst[t] = h;
                    }
                }
                out_base += wtot;
                total += wtot;
            }
        }
    }
    if (!FILL && lane == 0) counts[q] = (uint32_t)min(total, (uint64_t)0xFFFFFFFFu);
}

int query_run(pdg_graph *g, const uint64_t *h_a, const uint64_t *h_b, uint64_t n_q, int64_t max_len,
              bool connected_only, uint64_t *total) {
    cudaStream_t s = g->stream;
    if (!g->finalized) {
        set_error("query before pdg_finalize");
        return PDG_E_INVALID;
    }
    g->q_n = n_q;
    g->q_total = 0;
    *total = 0;
    g->q_off.reserve((n_q + 1) * 8, s);
    if (n_q == 0) {
        PDG_CUDA(cudaMemsetAsync(g->q_off.p, 0, 8, s));
        return PDG_OK;
    }
    g->q_a.reserve(n_q * 8, s);
    g->q_b.reserve(n_q * 8, s);
    g->q_cnt.reserve(n_q * 4, s);
    PDG_CUDA(cudaMemcpyAsync(g->q_a.p, h_a, n_q * 8, cudaMemcpyHostToDevice, s));
    PDG_CUDA(cudaMemcpyAsync(g->q_b.p, h_b, n_q * 8, cudaMemcpyHostToDevice, s));
    const GraphView G = view_of(g);
    const unsigned grid = (unsigned)((n_q * 32 + 255) / 256);
    query_kernel<false><<<grid, 256, 0, s>>>(G, g->q_a.as<uint64_t>(), g->q_b.as<uint64_t>(), n_q, max_len,
                                             connected_only ? 1 : 0, g->q_cnt.as<uint32_t>(), nullptr, nullptr);
    PDG_LAUNCH_CHECK();
    exclusive_scan_u32_u64(g->q_cnt.as<uint32_t>(), g->q_off.as<uint64_t>(), n_q, g);
    uint64_t tot = 0;
    PDG_CUDA(cudaMemcpyAsync(&tot, g->q_off.as<uint64_t>() + n_q, 8, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    g->q_total = tot;
    *total = tot;
    if (tot) {
        g->q_hits.reserve(tot * sizeof(pdg_hit), s);
        query_kernel<true><<<grid, 256, 0, s>>>(G, g->q_a.as<uint64_t>(), g->q_b.as<uint64_t>(), n_q, max_len,
                                                connected_only ? 1 : 0, nullptr, g->q_off.as<uint64_t>(),
                                                g->q_hits.as<pdg_hit>());
        PDG_LAUNCH_CHECK();
    }
    return PDG_OK;
}

// Follows the chain of edges (type, i), (type, i+1) ... (type, j) from the start edge's target
// (lemon_graph.py:244-264).  status: 0 ok, 1 chain broken, 2 more than one edge (type, m) leaves
// a node on the chain (the reference's `assert len(chains) == 1` would fire).
__global__ void walk_kernel(GraphView G, uint64_t src_code, pdg_hit hit, uint32_t *__restrict__ ids,
                            int32_t *__restrict__ status) {
    if (threadIdx.x || blockIdx.x) return;
    int32_t st = 0;
    const int64_t ia = find_node(G, src_code);
    if (ia < 0) { *status = 1; return; }
    ids[0] = (uint32_t)ia;
    uint32_t cur = hit.src_tgt_id;
    ids[1] = cur;
    uint32_t n = 2;
    for (int64_t m = (int64_t)hit.i + 1; m <= (int64_t)hit.j; ++m) {
        const uint64_t r0 = G.row_ptr[cur], r1 = G.row_ptr[cur + 1];
        const uint64_t key = ((uint64_t)hit.type_id << 32) | (uint64_t)((uint32_t)(int32_t)m ^ 0x80000000u);
        const uint64_t e = lower_bound_tv(G.csr_type, G.csr_val, r0, r1, key);
        if (e >= r1 || tv_key(G.csr_type, G.csr_val, e) != key) { st = 1; break; }
        if (e + 1 < r1 && tv_key(G.csr_type, G.csr_val, e + 1) == key) st = 2;
        cur = G.csr_tgt[e];
        ids[n++] = cur;
    }
    *status = st;
}

int query_walk(pdg_graph *g, uint64_t src_code, const pdg_hit *hit, uint32_t *h_ids, int32_t *status) {
    cudaStream_t s = g->stream;
    if (!g->finalized) {
        set_error("query before pdg_finalize");
        return PDG_E_INVALID;
    }
    if (hit->j < hit->i) {
        set_error("pdg_query_walk: j < i");
        return PDG_E_INVALID;
    }
    const uint64_t n = (uint64_t)((int64_t)hit->j - (int64_t)hit->i) + 2;
    g->small.reserve(n * 4 + 16, s);
    uint32_t *d_ids = g->small.as<uint32_t>() + 4;
    int32_t *d_st = g->small.as<int32_t>();
    PDG_CUDA(cudaMemsetAsync(g->small.p, 0, n * 4 + 16, s));
    walk_kernel<<<1, 32, 0, s>>>(view_of(g), src_code, *hit, d_ids, d_st);
    PDG_LAUNCH_CHECK();
    PDG_CUDA(cudaMemcpyAsync(h_ids, d_ids, n * 4, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaMemcpyAsync(status, d_st, 4, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    return PDG_OK;
}

}  // namespace pdg
