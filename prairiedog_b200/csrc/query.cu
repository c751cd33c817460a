// query.cu -- K4: batched "all paths between two k-mers".
//
// Restates LGGraph.connected / path (prairiedog/lemon_graph.py:178-334) and Graph.matching_edges
// (prairiedog/graph.py:82-113): a start edge (T,i) out of a matches every edge (T,j) into b with
// i <= j (and j - i <= max_len for path()).
//   genome part: a's row and b's row of the occurrence index are ascending position lists.  An
//     occurrence p of a inside contig c is a start edge (T=type[c], i=p-begin[c]) unless it is the
//     last k-mer of c; the matching in-edges of b are its occurrences q in [p+1, min(p+1+max_len,
//     end[c]-k)], one contiguous range of b's row found by binary search (j = q-begin[c]-1).
//   explicit part: CSR row of a and CSC row of b are sorted by (type, value); same range search.
// One warp per query; a count launch, a prefix sum, then a fill launch.
#include "gindex.cuh"

namespace pdg {

struct XView {
    const uint64_t *codes;
    uint64_t n_nodes;
    const uint64_t *row_ptr;
    const uint32_t *csr_tgt, *csr_type, *csr_val;
    const uint64_t *col_ptr;
    const uint32_t *csc_src, *csc_type, *csc_val;
};

struct GView {
    NodeLookup L;
    const uint32_t *rowptr, *pos;
    ContigTable ct;
    int k;
};

static XView xview_of(const pdg_graph *g) {
    XView v;
    v.codes = g->x_codes.as<uint64_t>();
    v.n_nodes = g->x_n_nodes;
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

static GView gview_of(const pdg_graph *g) {
    GView v;
    v.L = g_lookup(g);
    v.rowptr = g->g_rowptr.as<uint32_t>();
    v.pos = g->g_pos.as<uint32_t>();
    v.ct = g_table(g);
    v.k = g->k;
    return v;
}

__device__ __forceinline__ uint64_t warp_excl_scan(uint64_t cnt, int lane, uint64_t *wtot) {
    uint64_t inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint64_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    *wtot = __shfl_sync(0xffffffffu, inc, 31);
    return inc - cnt;
}

// first index in [lo,hi) with pos[index] >= key
__device__ __forceinline__ uint64_t lower_bound_pos(const uint32_t *pos, uint64_t lo, uint64_t hi, uint64_t key) {
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if ((uint64_t)pos[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// ---- genome part ----
template <bool FILL>
__global__ void __launch_bounds__(256) gquery_kernel(GView G, const uint64_t *__restrict__ qa,
                                                     const uint64_t *__restrict__ qb, uint64_t n_q, int64_t max_len,
                                                     int connected, uint32_t *__restrict__ counts,
                                                     const uint64_t *__restrict__ offsets, pdg_hit *__restrict__ hits) {
    const uint64_t q = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (q >= n_q) return;
    const int64_t ia = find_code(G.L, qa[q]);
    const int64_t ib = find_code(G.L, qb[q]);
    uint64_t total = 0;
    if (ia >= 0 && ib >= 0) {
        const uint64_t o0 = G.rowptr[ia], o1 = G.rowptr[ia + 1];
        const uint64_t i0 = G.rowptr[ib], i1 = G.rowptr[ib + 1];
        uint64_t out_base = FILL ? offsets[q] : 0;
        for (uint64_t base = o0; base < o1; base += 32) {
            const uint64_t e = base + lane;
            uint64_t lo = 0, cnt = 0, c = 0, p = 0, cb = 0;
            if (e < o1) {
                p = G.pos[e];
                c = contig_of(G.ct, p);
                cb = G.ct.begin[c];
                const uint64_t ce = G.ct.end[c];
                if (p + G.k < ce) {                       // not the last k-mer: an out-edge (T, p-cb)
                    uint64_t qmax = ce - G.k;            // last k-mer start of the contig
                    if (!connected && max_len >= 0) qmax = min(qmax, p + 1 + (uint64_t)max_len);
                    lo = lower_bound_pos(G.pos, i0, i1, p + 1);
                    const uint64_t hi = lower_bound_pos(G.pos, lo, i1, qmax + 1);
                    cnt = hi - lo;
                    if (connected) cnt = cnt ? 1 : 0;
                }
            }
            uint64_t wtot;
            const uint64_t excl = warp_excl_scan(cnt, lane, &wtot);
            if (FILL && cnt) {
                pdg_hit *dst = hits + out_base + excl;
                for (uint64_t t = 0; t < cnt; ++t) {
                    pdg_hit h;
                    h.type_id = G.ct.type[c];
                    h.i = (int32_t)(p - cb);
                    h.j = connected ? h.i : (int32_t)((uint64_t)G.pos[lo + t] - cb - 1);
                    h.aux = (uint32_t)c;
                    dst[t] = h;
                }
            }
            out_base += wtot;
            total += wtot;
        }
    }
    if (!FILL && lane == 0) counts[q] = (uint32_t)min(total, (uint64_t)0xFFFFFFFFu);
}

// ---- explicit part ----
__device__ __forceinline__ uint64_t tv_key(const uint32_t *type, const uint32_t *val, uint64_t e) {
    return ((uint64_t)type[e] << 32) | (uint64_t)val[e];
}
__device__ __forceinline__ uint64_t lower_bound_tv(const uint32_t *type, const uint32_t *val, uint64_t lo, uint64_t hi,
                                                   uint64_t key) {
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (tv_key(type, val, mid) < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}
__device__ __forceinline__ uint64_t upper_bound_tv(const uint32_t *type, const uint32_t *val, uint64_t lo, uint64_t hi,
                                                   uint64_t key) {
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (tv_key(type, val, mid) <= key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// counts are ADDED to counts[q]; hits are written after the genome-part hits (offsets[q] + gcount[q])
template <bool FILL>
__global__ void __launch_bounds__(256) xquery_kernel(XView G, const uint64_t *__restrict__ qa,
                                                     const uint64_t *__restrict__ qb, uint64_t n_q, int64_t max_len,
                                                     int connected, uint32_t *__restrict__ counts,
                                                     const uint32_t *__restrict__ gcount,
                                                     const uint64_t *__restrict__ offsets, pdg_hit *__restrict__ hits) {
    const uint64_t q = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (q >= n_q) return;
    const NodeLookup L{nullptr, nullptr, G.codes, G.n_nodes, 0};
    const int64_t ia = find_code(L, qa[q]);
    const int64_t ib = find_code(L, qb[q]);
    uint64_t total = 0;
    if (ia >= 0 && ib >= 0) {
        const uint64_t o0 = G.row_ptr[ia], o1 = G.row_ptr[ia + 1];
        const uint64_t i0 = G.col_ptr[ib], i1 = G.col_ptr[ib + 1];
        uint64_t out_base = FILL ? offsets[q] + gcount[q] : 0;
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
                uint64_t wtot;
                const uint64_t excl = warp_excl_scan(cnt, lane, &wtot);
                if (FILL && cnt) {
                    pdg_hit *dst = hits + out_base + excl;
                    for (uint64_t t = 0; t < cnt; ++t) {
                        pdg_hit h;
                        h.type_id = ty;
                        h.i = (int32_t)(vb ^ 0x80000000u);
                        h.j = connected ? h.i : (int32_t)(G.csc_val[lo + t] ^ 0x80000000u);
                        h.aux = PDG_HIT_EXPLICIT | G.csr_tgt[e];
                        dst[t] = h;
                    }
                }
                out_base += wtot;
                total += wtot;
            }
        }
    }
    if (!FILL && lane == 0) counts[q] = (uint32_t)min((uint64_t)counts[q] + total, (uint64_t)0xFFFFFFFFu);
}

__global__ void q_saturated_kernel(const uint32_t *__restrict__ counts, uint64_t n_q, uint32_t *__restrict__ n_sat) {
    const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q < n_q && counts[q] == 0xFFFFFFFFu) atomicAdd(n_sat, 1u);
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
    g->qs_total = 0;
    g->qs_ready = false;
    g->q_connected = connected_only;
    *total = 0;
    g->q_off.reserve((n_q + 1) * 8, s);
    if (n_q == 0) {
        PDG_CUDA(cudaMemsetAsync(g->q_off.p, 0, 8, s));
        return PDG_OK;
    }
    g->q_a.reserve(n_q * 8, s);
    g->q_b.reserve(n_q * 8, s);
    g->q_cnt.reserve(n_q * 4, s);
    g->q_cnt_g.reserve(n_q * 4, s);
    PDG_CUDA(cudaMemcpyAsync(g->q_a.p, h_a, n_q * 8, cudaMemcpyHostToDevice, s));
    PDG_CUDA(cudaMemcpyAsync(g->q_b.p, h_b, n_q * 8, cudaMemcpyHostToDevice, s));
    const uint64_t *qa = g->q_a.as<uint64_t>(), *qb = g->q_b.as<uint64_t>();
    uint32_t *cnt = g->q_cnt.as<uint32_t>(), *cnt_g = g->q_cnt_g.as<uint32_t>();
    const int conn = connected_only ? 1 : 0;
    const bool has_g = g->g_kmers != 0, has_x = g->x_n_edges != 0;
    const XView X = xview_of(g);
    const GView G = gview_of(g);
    const unsigned grid = (unsigned)((n_q * 32 + 255) / 256);
    PDG_CUDA(cudaMemsetAsync(cnt, 0, n_q * 4, s));
    if (has_g) {
        gquery_kernel<false><<<grid, 256, 0, s>>>(G, qa, qb, n_q, max_len, conn, cnt, nullptr, nullptr);
        PDG_LAUNCHED(g);
    }
    PDG_CUDA(cudaMemcpyAsync(cnt_g, cnt, n_q * 4, cudaMemcpyDeviceToDevice, s));
    if (has_x) {
        xquery_kernel<false><<<grid, 256, 0, s>>>(X, qa, qb, n_q, max_len, conn, cnt, cnt_g, nullptr, nullptr);
        PDG_LAUNCHED(g);
    }
    // the count kernels saturate counts[q] at 2^32 - 1 while the fill kernels write the TRUE number of hits:
    // a saturated count would let one query run into the next one's slots, so it is an error here
    g->small.reserve(16, s);
    PDG_CUDA(cudaMemsetAsync(g->small.p, 0, 4, s));
    q_saturated_kernel<<<(unsigned)((n_q + 255) / 256), 256, 0, s>>>(cnt, n_q, g->small.as<uint32_t>());
    PDG_LAUNCHED(g);
    exclusive_scan_u32_u64(cnt, g->q_off.as<uint64_t>(), n_q, g);
    uint64_t tot = 0;
    uint32_t saturated = 0;
    PDG_CUDA(cudaMemcpyAsync(&tot, g->q_off.as<uint64_t>() + n_q, 8, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaMemcpyAsync(&saturated, g->small.p, 4, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    if (saturated) {
        g->q_n = 0;
        set_error("query: %u quer%s more than 4294967294 hits each (32-bit hit counts); lower max_len or split the batch",
                  saturated, saturated == 1 ? "y has" : "ies have");
        return PDG_E_CAPACITY;
    }
    g->q_total = tot;
    *total = tot;
    if (tot) {
        g->q_hits.reserve(tot * sizeof(pdg_hit), s);
        if (has_g) {
            gquery_kernel<true><<<grid, 256, 0, s>>>(G, qa, qb, n_q, max_len, conn, nullptr, g->q_off.as<uint64_t>(),
                                                     g->q_hits.as<pdg_hit>());
            PDG_LAUNCHED(g);
        }
        if (has_x) {
            xquery_kernel<true><<<grid, 256, 0, s>>>(X, qa, qb, n_q, max_len, conn, nullptr, cnt_g,
                                                     g->q_off.as<uint64_t>(), g->q_hits.as<pdg_hit>());
            PDG_LAUNCHED(g);
        }
    }
    return PDG_OK;
}

// Explicit part: follows the chain of edges (type, i), (type, i+1) ... (type, j) from the start
// edge's target (lemon_graph.py:244-264).  status: 0 ok, 1 chain broken, 2 more than one edge
// (type, m) leaves a node on the chain (the reference's `assert len(chains) == 1` would fire).
__global__ void xwalk_kernel(XView G, uint64_t src_code, pdg_hit hit, uint64_t *__restrict__ codes,
                             int32_t *__restrict__ status) {
    if (threadIdx.x || blockIdx.x) return;
    int32_t st = 0;
    codes[0] = src_code;
    uint32_t cur = hit.aux & ~PDG_HIT_EXPLICIT;
    codes[1] = G.codes[cur];
    uint32_t n = 2;
    for (int64_t m = (int64_t)hit.i + 1; m <= (int64_t)hit.j; ++m) {
        const uint64_t r0 = G.row_ptr[cur], r1 = G.row_ptr[cur + 1];
        const uint64_t key = ((uint64_t)hit.type_id << 32) | (uint64_t)((uint32_t)(int32_t)m ^ 0x80000000u);
        const uint64_t e = lower_bound_tv(G.csr_type, G.csr_val, r0, r1, key);
        if (e >= r1 || tv_key(G.csr_type, G.csr_val, e) != key) { st = 1; break; }
        if (e + 1 < r1 && tv_key(G.csr_type, G.csr_val, e + 1) == key) st = 2;
        cur = G.csr_tgt[e];
        codes[n++] = G.codes[cur];
    }
    *status = st;
}

int query_walk(pdg_graph *g, uint64_t src_code, const pdg_hit *hit, uint64_t *h_codes, int32_t *status) {
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
    g->small.reserve(n * 8 + 16, s);
    uint64_t *d_codes = g->small.as<uint64_t>() + 2;
    int32_t *d_st = g->small.as<int32_t>();
    PDG_CUDA(cudaMemsetAsync(g->small.p, 0, n * 8 + 16, s));
    if (hit->aux & PDG_HIT_EXPLICIT) {
        if ((hit->aux & ~PDG_HIT_EXPLICIT) >= g->x_n_nodes) {
            set_error("pdg_query_walk: bad hit");
            return PDG_E_INVALID;
        }
        xwalk_kernel<<<1, 32, 0, s>>>(xview_of(g), src_code, *hit, d_codes, d_st);
        PDG_LAUNCHED(g);
    } else {
        // genome part: the path is the contig slice starting at begin + i (lemon_graph.py:244-264
        // follows exactly these edges because (type, value) is unique inside G)
        if (hit->aux >= g->g_begin.size() || hit->i < 0 ||
            (uint64_t)hit->j + 1 + g->k >= g->g_len[hit->aux] + 1) {
            set_error("pdg_query_walk: bad hit");
            return PDG_E_INVALID;
        }
        const int rc = emit_codes(g, g->g_begin[hit->aux] + (uint64_t)hit->i, n + g->k - 1, d_codes);
        if (rc != PDG_OK) return rc;
    }
    PDG_CUDA(cudaMemcpyAsync(h_codes, d_codes, n * 8, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaMemcpyAsync(status, d_st, 4, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    return PDG_OK;
}

// ---- strings of a batch's hits ---------------------------------------------------------------------------
// Prairiedog.query (prairiedog/prairiedog.py:22-31) spells every hit with concat_values (prairiedog/node.py:22-37):
// first node value + last character of every later one = residues [begin + i, begin + j + 1 + k) of the hit's
// contig.  One pass sizes the strings (16-byte aligned starts so that a warp stores whole uint4), one pass decodes
// them straight off the 2-bit store: a warp per hit, a lane per 16 characters, 512 contiguous bytes per warp store.
// Hits of the explicit part need the chain walk of lemon_graph.py:244-264 (node values can be opaque there): they
// get length 0 here and the host spells them with pdg_query_walk.
__global__ void qs_len_kernel(const pdg_hit *__restrict__ hits, uint64_t n, int k, uint32_t *__restrict__ len16) {
    const uint64_t h = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= n) return;
    const pdg_hit x = hits[h];
    const uint32_t len = (x.aux & PDG_HIT_EXPLICIT) ? 0u : (uint32_t)(x.j - x.i + 1 + k);
    len16[h] = (len + 15u) >> 4;
}

__global__ void __launch_bounds__(256) qs_fill_kernel(const pdg_hit *__restrict__ hits, const uint64_t *__restrict__ off16,
                                                      uint64_t n, const uint64_t *__restrict__ cbegin,
                                                      const unsigned long long *__restrict__ packed, int k,
                                                      uint4 *__restrict__ out) {
    const uint64_t h = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (h >= n) return;
    const pdg_hit x = hits[h];
    if (x.aux & PDG_HIT_EXPLICIT) return;
    const uint64_t p0 = cbegin[x.aux] + (uint64_t)x.i;
    const uint32_t len = (uint32_t)(x.j - x.i + 1 + k);
    uint4 *dst = out + off16[h];
    for (uint32_t c = 16u * lane; c < len; c += 512u) {
        // 16 residues from c on as one 32-bit code (reads past the string stay inside the zero-padded store)
        const uint32_t code = (uint32_t)kmer_at(packed, p0 + c, 16);
        uint32_t w[4];
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            const uint32_t b = (code >> (24 - 8 * m)) & 0xffu;    // residues 4m .. 4m+3, first one on top
            const uint32_t sel = (b >> 6) | (((b >> 4) & 3u) << 4) | (((b >> 2) & 3u) << 8) | ((b & 3u) << 12);
            w[m] = __byte_perm(0x47435441u, 0u, sel);             // "ATCG"[digit], one byte per residue
            const uint32_t left = len - c > 4u * m ? len - c - 4u * m : 0u;   // characters of the string in this word
            if (left < 4u) w[m] &= left ? (0xffffffffu >> (32u - 8u * left)) : 0u;
        }
        dst[c >> 4] = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// sizes the strings of the hits of the last pdg_query_paths call; *total_bytes = size of the byte buffer
int query_strings_size(pdg_graph *g, uint64_t *total_bytes) {
    cudaStream_t s = g->stream;
    if (!g->finalized) {
        set_error("pdg_query_strings_size before pdg_finalize");
        return PDG_E_INVALID;
    }
    if (g->q_connected) {
        set_error("pdg_query_strings_size: the last query was pdg_query_connected (start edges, not paths)");
        return PDG_E_INVALID;
    }
    const uint64_t n = g->q_total;
    g->qs_off.reserve((n + 1) * 8, s);
    g->qs_total = 0;
    if (n == 0) {
        PDG_CUDA(cudaMemsetAsync(g->qs_off.p, 0, 8, s));
    } else {
        g->qs_len.reserve(n * 4, s);
        qs_len_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(g->q_hits.as<pdg_hit>(), n, g->k, g->qs_len.as<uint32_t>());
        PDG_LAUNCHED(g);
        exclusive_scan_u32_u64(g->qs_len.as<uint32_t>(), g->qs_off.as<uint64_t>(), n, g);
        uint64_t tot16 = 0;
        PDG_CUDA(cudaMemcpyAsync(&tot16, g->qs_off.as<uint64_t>() + n, 8, cudaMemcpyDeviceToHost, s));
        PDG_CUDA(cudaStreamSynchronize(s));
        g->qs_total = tot16 * 16;
    }
    g->qs_ready = true;
    *total_bytes = g->qs_total;
    return PDG_OK;
}

// h_offsets[n_hits + 1]: byte offset of every hit's string (its length is j - i + 1 + k; 0 for explicit hits);
// h_bytes: qs_total bytes.  h_bytes may be pinned memory (one cudaMemcpyAsync straight into it).
int query_strings_fetch(pdg_graph *g, uint64_t *h_offsets, uint8_t *h_bytes) {
    cudaStream_t s = g->stream;
    if (!g->qs_ready) {
        set_error("pdg_query_strings_fetch before pdg_query_strings_size");
        return PDG_E_INVALID;
    }
    const uint64_t n = g->q_total;
    if (g->qs_total) {
        g->qs_bytes.reserve(g->qs_total, s);
        const ContigTable ct = g_table(g);
        qs_fill_kernel<<<(unsigned)((n * 32 + 255) / 256), 256, 0, s>>>(g->q_hits.as<pdg_hit>(), g->qs_off.as<uint64_t>(), n,
                                                                       ct.begin, g->packed.as<unsigned long long>(), g->k,
                                                                       g->qs_bytes.as<uint4>());
        PDG_LAUNCHED(g);
        if (h_bytes) PDG_CUDA(cudaMemcpyAsync(h_bytes, g->qs_bytes.p, g->qs_total, cudaMemcpyDeviceToHost, s));
    }
    if (h_offsets) {
        PDG_CUDA(cudaMemcpyAsync(h_offsets, g->qs_off.p, (n + 1) * 8, cudaMemcpyDeviceToHost, s));
        PDG_CUDA(cudaStreamSynchronize(s));
        for (uint64_t i = 0; i <= n; ++i) h_offsets[i] *= 16;   // the device keeps them in 16-byte units
    }
    PDG_CUDA(cudaStreamSynchronize(s));
    return PDG_OK;
}

}  // namespace pdg
