// xbuild.cu -- the explicit part: edges / nodes given through the per-edge API (and contigs that
// had to be expanded because their type id is shared) -> deduplicated CSR + CSC.
//   prairiedog/lemon_graph.py:99-128      LGGraph.upsert_node / add_edge
//   lemongraph/lib/lemongraph.c:661-672   node upsert unique on (type,value) -> sort + run heads
//   lemongraph/lib/lemongraph.c:674-685   edge upsert unique on (type,value,src,tgt) -> sort + run heads
//   lemongraph/lib/lemongraph.c:603-630   srcnode_idx / tgtnode_idx -> CSR / CSC
#include "gindex.cuh"
#include <algorithm>

namespace pdg {

__global__ void head_flags_u64_kernel(const uint64_t *__restrict__ keys, uint64_t n, uint32_t *__restrict__ flags) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1u : 0u;
}
__global__ void compact_u64_kernel(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ flags,
                                   const uint64_t *__restrict__ pos, uint64_t n, uint64_t *__restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && flags[i]) out[pos[i]] = keys[i];
}
__device__ __forceinline__ bool rec_eq(const EdgeRec &a, const EdgeRec &b) {
    return a.x == b.x && a.y == b.y && a.z == b.z && a.w == b.w;
}
__global__ void head_flags_rec_kernel(const EdgeRec *__restrict__ r, uint64_t n, uint32_t *__restrict__ flags) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = (i == 0 || !rec_eq(r[i], r[i - 1])) ? 1u : 0u;
}
__global__ void compact_rec_kernel(const EdgeRec *__restrict__ r, const uint32_t *__restrict__ flags,
                                   const uint64_t *__restrict__ pos, uint64_t n, EdgeRec *__restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && flags[i]) out[pos[i]] = r[i];
}

// (src code, tgt code, type, value) -> record over X node ids
__global__ void gen_explicit_kernel(const uint64_t *__restrict__ src, const uint64_t *__restrict__ tgt,
                                    const uint32_t *__restrict__ type, const int32_t *__restrict__ val, uint64_t n,
                                    EdgeRec *__restrict__ out, NodeLookup L) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    EdgeRec r;
    r.x = (uint32_t)find_code(L, src[i]);
    r.y = (uint32_t)find_code(L, tgt[i]);
    r.z = type[i];
    r.w = (uint32_t)val[i] ^ 0x80000000u;
    out[i] = r;
}

// sorted records -> row pointers + structure-of-arrays payload
// KEYF = 0: CSR (key = src id x, other = tgt id y);  KEYF = 1: CSC (key = tgt id y, other = src id x)
template <int KEYF>
__global__ void pack_rows_kernel(const EdgeRec *__restrict__ r, uint64_t n, uint64_t n_nodes,
                                 uint64_t *__restrict__ ptr, uint32_t *__restrict__ other,
                                 uint32_t *__restrict__ type, uint32_t *__restrict__ val) {
    const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    const EdgeRec x = r[e];
    const uint32_t key = KEYF == 0 ? x.x : x.y;
    other[e] = KEYF == 0 ? x.y : x.x;
    type[e] = x.z;
    val[e] = x.w;
    if (e == 0) {
        for (uint64_t id = 0; id <= key; ++id) ptr[id] = 0;
    } else {
        const EdgeRec pr = r[e - 1];
        const uint32_t pk = KEYF == 0 ? pr.x : pr.y;
        for (uint64_t id = (uint64_t)pk + 1; id <= key; ++id) ptr[id] = e;
    }
    if (e == n - 1)
        for (uint64_t id = (uint64_t)key + 1; id <= n_nodes; ++id) ptr[id] = n;
}

uint64_t sort_unique_keys(pdg_graph *g, uint64_t *keys, uint64_t *tmp, uint64_t n, int key_bits, DevBuf &out) {
    cudaStream_t s = g->stream;
    out.reserve(16, s);
    if (n == 0) return 0;
    uint64_t pr = 0;
    const uint64_t *sorted = radix_sort_keys(g, keys, tmp, n, 0, key_bits, &pr);
    g->scratch0.reserve(n * 4, s);
    g->scratch1.reserve((n + 1) * 8, s);
    uint32_t *flg = g->scratch0.as<uint32_t>();
    uint64_t *posn = g->scratch1.as<uint64_t>();
    head_flags_u64_kernel<<<grid_for(n, 256), 256, 0, s>>>(sorted, n, flg);
    PDG_LAUNCHED(g);
    exclusive_scan_u32_u64(flg, posn, n, g);
    uint64_t nn = 0;
    PDG_CUDA(cudaMemcpyAsync(&nn, posn + n, 8, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    out.reserve(std::max<size_t>(nn * 8, 16), s);
    compact_u64_kernel<<<grid_for(n, 256), 256, 0, s>>>(sorted, flg, posn, n, out.as<uint64_t>());
    PDG_LAUNCHED(g);
    return nn;
}

static std::vector<PassDesc> field_passes(int field, int bits) {
    std::vector<PassDesc> v;
    for (int s = 0; s < bits; s += 8) v.push_back(PassDesc{field, s});
    return v;
}

void build_x_part(pdg_graph *g, const HostContigs &dem) {
    cudaStream_t s = g->stream;
    const uint64_t n_api = g->x_src.size(), n_dem = dem.n_kmers - dem.size(), n_rec = n_api + n_dem;
    const uint64_t n_up = g->x_nodes.size();
    g->x_n_nodes = g->x_n_edges = 0;
    g->row_ptr.reserve(16, s);
    g->col_ptr.reserve(16, s);
    g->x_codes.reserve(16, s);
    PDG_CUDA(cudaMemsetAsync(g->row_ptr.p, 0, 16, s));
    PDG_CUDA(cudaMemsetAsync(g->col_ptr.p, 0, 16, s));
    if (n_rec == 0 && n_up == 0) return;

    // ---- explicit records (src code, tgt code, type, value) on the device ----
    DevBuf d_src, d_tgt, d_type, d_val, d_keys, d_tmp;
    d_src.reserve(std::max<size_t>(n_rec * 8, 16), s);
    d_tgt.reserve(std::max<size_t>(n_rec * 8, 16), s);
    d_type.reserve(std::max<size_t>(n_rec * 4, 16), s);
    d_val.reserve(std::max<size_t>(n_rec * 4, 16), s);
    if (n_api) {
        PDG_CUDA(cudaMemcpyAsync(d_src.p, g->x_src.data(), n_api * 8, cudaMemcpyHostToDevice, s));
        PDG_CUDA(cudaMemcpyAsync(d_tgt.p, g->x_tgt.data(), n_api * 8, cudaMemcpyHostToDevice, s));
        PDG_CUDA(cudaMemcpyAsync(d_type.p, g->x_type.data(), n_api * 4, cudaMemcpyHostToDevice, s));
        PDG_CUDA(cudaMemcpyAsync(d_val.p, g->x_val.data(), n_api * 4, cudaMemcpyHostToDevice, s));
    }
    if (n_dem) {
        DevBuf db, de, dk, dt;
        const ContigTable ct = dem.upload(db, de, dk, dt, s);
        TraverseOut out;
        out.x_src = d_src.as<uint64_t>() + n_api;
        out.x_tgt = d_tgt.as<uint64_t>() + n_api;
        out.x_type = d_type.as<uint32_t>() + n_api;
        out.x_val = d_val.as<int32_t>() + n_api;
        run_traverse(g, MODE_XGEN, ct, dem.pos0, dem.pos1, out);
        PDG_CUDA(cudaStreamSynchronize(s));   // db..dt die with this scope
    }
    // ---- X node table: endpoints + upserted nodes ----
    const uint64_t n_keys = 2 * n_rec + n_up;
    d_keys.reserve(n_keys * 8, s);
    d_tmp.reserve(n_keys * 8, s);
    if (n_rec) {
        PDG_CUDA(cudaMemcpyAsync(d_keys.p, d_src.p, n_rec * 8, cudaMemcpyDeviceToDevice, s));
        PDG_CUDA(cudaMemcpyAsync(d_keys.as<uint64_t>() + n_rec, d_tgt.p, n_rec * 8, cudaMemcpyDeviceToDevice, s));
    }
    if (n_up)
        PDG_CUDA(cudaMemcpyAsync(d_keys.as<uint64_t>() + 2 * n_rec, g->x_nodes.data(), n_up * 8, cudaMemcpyHostToDevice, s));
    g->x_n_nodes = sort_unique_keys(g, d_keys.as<uint64_t>(), d_tmp.as<uint64_t>(), n_keys, 64, g->x_codes);
    if (g->x_n_nodes >= (1ULL << 31)) {
        set_error("pdg_finalize: %llu explicit nodes exceed 31-bit ids", (unsigned long long)g->x_n_nodes);
        throw CudaError{PDG_E_INVALID};
    }
    const uint64_t nn = g->x_n_nodes;
    g->row_ptr.reserve((nn + 1) * 8, s);
    g->col_ptr.reserve((nn + 1) * 8, s);
    PDG_CUDA(cudaMemsetAsync(g->row_ptr.p, 0, (nn + 1) * 8, s));
    PDG_CUDA(cudaMemsetAsync(g->col_ptr.p, 0, (nn + 1) * 8, s));
    if (n_rec == 0) return;

    // ---- records over node ids, full-key sort, dedup, CSR; re-sort by target, CSC ----
    NodeLookup L{nullptr, nullptr, g->x_codes.as<uint64_t>(), nn, 0};
    DevBuf ra_b, rb_b;
    ra_b.reserve(n_rec * sizeof(EdgeRec), s);
    rb_b.reserve(n_rec * sizeof(EdgeRec), s);
    EdgeRec *ra = ra_b.as<EdgeRec>(), *rb = rb_b.as<EdgeRec>();
    gen_explicit_kernel<<<grid_for(n_rec, 256), 256, 0, s>>>(d_src.as<uint64_t>(), d_tgt.as<uint64_t>(),
                                                            d_type.as<uint32_t>(), d_val.as<int32_t>(), n_rec, ra, L);
    PDG_LAUNCHED(g);
    const int id_bits = std::max(1, bit_length(nn ? nn - 1 : 0));
    auto key_passes = [&](int key_field) {   // least significant first: other endpoint, value, type, key
        std::vector<PassDesc> p = field_passes(key_field == 0 ? 1 : 0, id_bits);
        auto b = field_passes(3, 32), c = field_passes(2, 32), d = field_passes(key_field, id_bits);
        p.insert(p.end(), b.begin(), b.end());
        p.insert(p.end(), c.begin(), c.end());
        p.insert(p.end(), d.begin(), d.end());
        return p;
    };
    uint64_t pr = 0;
    EdgeRec *sorted = radix_sort_records(g, ra, rb, n_rec, key_passes(0), &pr);
    EdgeRec *other = sorted == ra ? rb : ra;
    g->scratch0.reserve(n_rec * 4, s);
    g->scratch1.reserve((n_rec + 1) * 8, s);
    uint32_t *flg = g->scratch0.as<uint32_t>();
    uint64_t *posn = g->scratch1.as<uint64_t>();
    head_flags_rec_kernel<<<grid_for(n_rec, 256), 256, 0, s>>>(sorted, n_rec, flg);
    PDG_LAUNCHED(g);
    exclusive_scan_u32_u64(flg, posn, n_rec, g);
    uint64_t n_e = 0;
    PDG_CUDA(cudaMemcpyAsync(&n_e, posn + n_rec, 8, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    compact_rec_kernel<<<grid_for(n_rec, 256), 256, 0, s>>>(sorted, flg, posn, n_rec, other);
    PDG_LAUNCHED(g);
    sorted = other;
    g->x_n_edges = n_e;
    for (DevBuf *b : {&g->csr_tgt, &g->csr_type, &g->csr_val, &g->csc_src, &g->csc_type, &g->csc_val})
        b->reserve(n_e * 4, s);
    pack_rows_kernel<0><<<grid_for(n_e, 256), 256, 0, s>>>(sorted, n_e, nn, g->row_ptr.as<uint64_t>(),
                                                          g->csr_tgt.as<uint32_t>(), g->csr_type.as<uint32_t>(),
                                                          g->csr_val.as<uint32_t>());
    PDG_LAUNCHED(g);
    EdgeRec *tmp2 = sorted == ra ? rb : ra;
    EdgeRec *sorted2 = radix_sort_records(g, sorted, tmp2, n_e, key_passes(1), &pr);
    pack_rows_kernel<1><<<grid_for(n_e, 256), 256, 0, s>>>(sorted2, n_e, nn, g->col_ptr.as<uint64_t>(),
                                                          g->csc_src.as<uint32_t>(), g->csc_type.as<uint32_t>(),
                                                          g->csc_val.as<uint32_t>());
    PDG_LAUNCHED(g);
    PDG_CUDA(cudaStreamSynchronize(s));   // local DevBufs are freed on return
}

}  // namespace pdg
