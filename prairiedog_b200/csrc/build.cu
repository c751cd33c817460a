// build.cu -- K1b/K2/K3: k-mer emission from the packed store, node dedup -> dense ids,
// labelled edge records -> CSR (by source) and CSC (by target).
//
// Restates, as bulk data-parallel passes, what the reference does one edge at a time:
//   prairiedog/kmers.py:96-124            sliding window  -> kmer_at()
//   prairiedog/subgraph_ref.py:62-102     edge (kmer_i -> kmer_i+1, type, value=i) -> gen_records
//   lemongraph/lib/lemongraph.c:661-672   node upsert unique on (type,value) -> bitmap / sort+RLE
//   lemongraph/lib/lemongraph.c:674-685   edge upsert unique on (type,value,src,tgt) -> sort+RLE
//   lemongraph/lib/lemongraph.c:603-630   srcnode_idx / tgtnode_idx -> CSR / CSC
#include "common.cuh"
#include <algorithm>
#include <chrono>

namespace pdg {

static constexpr int EM_THREADS = 256;
static constexpr int EM_ITEMS = 8;
static constexpr int EM_TILE = EM_THREADS * EM_ITEMS;      // residue positions per CTA
static constexpr int EM_WORDS = EM_TILE / 32 + 2;          // packed words staged per CTA

struct ContigTable {
    const uint64_t *begin;   // ascending
    const uint64_t *end;
    const uint64_t *ebase;   // edges before this contig (exclusive prefix of len-k)
    const uint32_t *type;
    uint64_t n;
};

struct NodeLookup {
    const uint32_t *bitmap;      // presence bits (bitmap path) or nullptr
    const uint32_t *prefix;      // exclusive popcount prefix per 32-bit word
    const uint64_t *codes;       // sorted unique codes (sort path)
    uint64_t n_nodes;
};

__device__ __forceinline__ uint32_t node_id(const NodeLookup &L, uint64_t code) {
    if (L.bitmap) {
        const uint32_t w = (uint32_t)(code >> 5), b = (uint32_t)code & 31u;
        return L.prefix[w] + __popc(L.bitmap[w] & ((1u << b) - 1u));
    }
    uint64_t lo = 0, hi = L.n_nodes;   // lower_bound
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (L.codes[mid] < code) lo = mid + 1; else hi = mid;
    }
    return (uint32_t)lo;
}

// k-mer starting at tile-local residue slot q of the staged words (big endian 2-bit packing)
__device__ __forceinline__ uint64_t kmer_at(const unsigned long long *words, uint32_t q, int k) {
    const uint32_t w = q >> 5, s = 2u * (q & 31u);
    const unsigned long long hi = words[w], lo = words[w + 1];
    const unsigned long long x = s ? ((hi << s) | (lo >> (64 - s))) : hi;
    return x >> (64 - 2 * k);
}

enum { MODE_MARK = 0, MODE_CODES = 1, MODE_GEN = 2 };

// One pass over residue positions [pos0, pos1).  Position p inside contig c (begin <= p,
// p + k <= end) is k-mer number p - begin of that contig; it is the source of edge
// (type[c], p - begin) when p + k < end.
template <int MODE>
__global__ void __launch_bounds__(EM_THREADS) traverse_kernel(const unsigned long long *__restrict__ packed,
                                                              uint64_t packed_words, uint64_t pos0,
                                                              uint64_t pos1, int k, ContigTable ct,
                                                              uint32_t *__restrict__ bitmap,      // MARK
                                                              uint64_t *__restrict__ codes_out,   // CODES
                                                              EdgeRec *__restrict__ rec_out,      // GEN
                                                              NodeLookup L) {
    __shared__ unsigned long long words[EM_WORDS];
    __shared__ uint64_t ci0_s;
    const uint64_t tile_p0 = pos0 + (uint64_t)blockIdx.x * EM_TILE;
    const uint64_t w0 = tile_p0 >> 5;
    if (threadIdx.x < EM_WORDS)
        words[threadIdx.x] = (w0 + threadIdx.x < packed_words) ? packed[w0 + threadIdx.x] : 0ULL;
    if (threadIdx.x == 0) {
        // last contig whose begin <= tile_p0 (or 0)
        uint64_t lo = 0, hi = ct.n;
        while (lo < hi) {
            const uint64_t mid = (lo + hi) >> 1;
            if (ct.begin[mid] <= tile_p0) lo = mid + 1; else hi = mid;
        }
        ci0_s = lo ? lo - 1 : 0;
    }
    __syncthreads();
    const uint32_t lead = (uint32_t)(tile_p0 & 31u);
    uint64_t ci = ci0_s;
#pragma unroll
    for (int j = 0; j < EM_ITEMS; ++j) {
        const uint32_t l = j * EM_THREADS + threadIdx.x;
        const uint64_t p = tile_p0 + l;
        if (p >= pos1) break;
        while (ci + 1 < ct.n && ct.begin[ci + 1] <= p) ++ci;
        const uint64_t cb = ct.begin[ci], ce = ct.end[ci];
        if (p < cb || p + k > ce) continue;
        const uint64_t code = kmer_at(words, lead + l, k);
        if (MODE == MODE_MARK) {
            const uint32_t w = (uint32_t)(code >> 5), bit = 1u << ((uint32_t)code & 31u);
            if (!(bitmap[w] & bit)) atomicOr(&bitmap[w], bit);
        } else if (MODE == MODE_CODES) {
            codes_out[ct.ebase[ci] + ci + (p - cb)] = code;   // k-mer ordinal = edges before + contigs before
        } else {
            if (p + k < ce) {
                const uint64_t mask = (k == 32) ? ~0ULL : ((1ULL << (2 * k)) - 1ULL);
                // next k-mer: drop the first base, append residue p+k
                const uint32_t q = lead + l + k;
                const uint64_t nb = (words[q >> 5] >> (62 - 2 * (q & 31u))) & 3ULL;
                const uint64_t tcode = ((code << 2) | nb) & mask;
                EdgeRec r;
                r.x = node_id(L, code);
                r.y = node_id(L, tcode);
                r.z = ct.type[ci];
                r.w = (uint32_t)(p - cb) ^ 0x80000000u;
                rec_out[ct.ebase[ci] + (p - cb)] = r;
            }
        }
    }
}

// ---- explicit (per-edge API) inputs -----------------------------------------------------------
__global__ void mark_codes_kernel(const uint64_t *__restrict__ codes, uint64_t n, uint32_t *__restrict__ bitmap) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t c = codes[i];
    atomicOr(&bitmap[c >> 5], 1u << ((uint32_t)c & 31u));
}

__global__ void gen_explicit_kernel(const uint64_t *__restrict__ src, const uint64_t *__restrict__ tgt,
                                    const uint32_t *__restrict__ type, const int32_t *__restrict__ val, uint64_t n,
                                    EdgeRec *__restrict__ out, NodeLookup L) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    EdgeRec r;
    r.x = node_id(L, src[i]);
    r.y = node_id(L, tgt[i]);
    r.z = type[i];
    r.w = (uint32_t)val[i] ^ 0x80000000u;
    out[i] = r;
}

// ---- bitmap -> ranks -> node table ---------------------------------------------------------------
__global__ void popc_words_kernel(const uint32_t *__restrict__ bm, uint64_t n_words, uint32_t *__restrict__ cnt) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_words) cnt[i] = __popc(bm[i]);
}

__global__ void bitmap_fill_codes_kernel(const uint32_t *__restrict__ bm, const uint32_t *__restrict__ prefix,
                                         uint64_t n_words, uint64_t *__restrict__ codes) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_words) return;
    uint32_t m = bm[i];
    uint32_t o = prefix[i];
    while (m) {
        const int b = __ffs(m) - 1;
        codes[o++] = (i << 5) | (uint64_t)b;
        m &= m - 1;
    }
}

// ---- run-length compaction helpers (sorted input) --------------------------------------------------
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

// ---- sorted records -> row pointers + structure-of-arrays payload ----------------------------------
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

// ---- checksums over the final structure ------------------------------------------------------------
__global__ void checksum_nodes_kernel(const uint64_t *__restrict__ codes, uint64_t n, unsigned long long *sum) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long h = i < n ? mix64(codes[i] + 0x9e3779b97f4a7c15ULL) : 0ULL;
    for (int o = 16; o; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
    if ((threadIdx.x & 31) == 0 && h) atomicAdd(sum, h);
}
// rows: ptr[n_nodes+1]; entry e of row `key`: (other, type, val).  src_is_row: CSR (row = src)
__global__ void checksum_edges_kernel(const uint64_t *__restrict__ ptr, uint64_t n_nodes, uint64_t n_edges,
                                      const uint32_t *__restrict__ other, const uint32_t *__restrict__ type,
                                      const uint32_t *__restrict__ val, const uint64_t *__restrict__ codes,
                                      int src_is_row, unsigned long long *sum) {
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
        h = mix64(s + 0x9e3779b97f4a7c15ULL);
        h = mix64(h ^ t);
        h = mix64(h ^ (((uint64_t)type[e] << 32) | (uint64_t)(val[e] ^ 0x80000000u)));
    }
    for (int o = 16; o; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
    if ((threadIdx.x & 31) == 0 && h) atomicAdd(sum, h);
}

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

static inline unsigned grid_for(uint64_t n, int threads) { return (unsigned)((n + threads - 1) / threads); }
static inline int bit_length(uint64_t v) { int b = 0; while (v) { ++b; v >>= 1; } return b; }

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

template <typename T>
static void upload_vec(DevBuf &buf, const std::vector<T> &v, cudaStream_t s) {
    buf.reserve(std::max<size_t>(v.size() * sizeof(T), 16), s);
    if (!v.empty()) PDG_CUDA(cudaMemcpyAsync(buf.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, s));
}

static ContigTable make_table(const std::vector<uint64_t> &begin, const std::vector<uint64_t> &len,
                              const std::vector<uint64_t> &ebase, const std::vector<uint32_t> &type, DevBuf &db,
                              DevBuf &de, DevBuf &dx, DevBuf &dt, cudaStream_t s) {
    std::vector<uint64_t> end(begin.size());
    for (size_t i = 0; i < begin.size(); ++i) end[i] = begin[i] + len[i];
    upload_vec(db, begin, s);
    upload_vec(de, end, s);
    upload_vec(dx, ebase, s);
    upload_vec(dt, type, s);
    PDG_CUDA(cudaStreamSynchronize(s));   // `end` is a temporary
    ContigTable ct;
    ct.begin = db.as<uint64_t>();
    ct.end = de.as<uint64_t>();
    ct.ebase = dx.as<uint64_t>();
    ct.type = dt.as<uint32_t>();
    ct.n = begin.size();
    return ct;
}

void upload_contigs(pdg_graph *g) {
    (void)make_table(g->c_begin, g->c_len, g->c_edge_base, g->c_type, g->d_cbegin, g->d_cend, g->d_cebase,
                     g->d_ctype, g->stream);
}

template <int MODE>
static void run_traverse(pdg_graph *g, const ContigTable &ct, uint64_t pos0, uint64_t pos1, uint32_t *bitmap,
                         uint64_t *codes, EdgeRec *recs, const NodeLookup &L) {
    if (ct.n == 0 || pos1 <= pos0) return;
    const uint64_t n_tiles = (pos1 - pos0 + EM_TILE - 1) / EM_TILE;
    traverse_kernel<MODE><<<(unsigned)n_tiles, EM_THREADS, 0, g->stream>>>(
        g->packed.as<unsigned long long>(), g->packed.cap / 8, pos0, pos1, g->k, ct, bitmap, codes, recs, L);
    PDG_LAUNCH_CHECK();
}

// span of residue positions covered by a contig table
static void table_span(const std::vector<uint64_t> &begin, const std::vector<uint64_t> &len, uint64_t *p0,
                       uint64_t *p1) {
    *p0 = begin.empty() ? 0 : begin.front();
    uint64_t e = 0;
    for (size_t i = 0; i < begin.size(); ++i) e = std::max(e, begin[i] + len[i]);
    *p1 = e;
}

static std::vector<PassDesc> field_passes(int field, int bits) {
    std::vector<PassDesc> v;
    for (int s = 0; s < bits; s += 8) v.push_back(PassDesc{field, s});
    return v;
}

int finalize(pdg_graph *g, int flags) {
    cudaStream_t s = g->stream;
    Timer total(s), lap(s);
    pdg_build_stats st{};
    g->stats = st;
    const int k = g->k;
    const uint64_t n_x = g->x_src.size();
    const uint64_t n_rec = g->n_gen_edges + n_x;

    // ---- contig table + explicit inputs to the device ----
    ContigTable ct = make_table(g->c_begin, g->c_len, g->c_edge_base, g->c_type, g->d_cbegin, g->d_cend,
                                g->d_cebase, g->d_ctype, s);
    uint64_t pos0, pos1;
    table_span(g->c_begin, g->c_len, &pos0, &pos1);
    // explicit node codes: endpoints of explicit edges + upserted nodes
    std::vector<uint64_t> xcodes;
    xcodes.reserve(2 * n_x + g->x_nodes.size());
    xcodes.insert(xcodes.end(), g->x_src.begin(), g->x_src.end());
    xcodes.insert(xcodes.end(), g->x_tgt.begin(), g->x_tgt.end());
    xcodes.insert(xcodes.end(), g->x_nodes.begin(), g->x_nodes.end());
    uint64_t max_x = 0;
    for (uint64_t c : xcodes) max_x = std::max(max_x, c);
    DevBuf d_xcodes, d_xsrc, d_xtgt, d_xtype, d_xval;
    upload_vec(d_xcodes, xcodes, s);
    upload_vec(d_xsrc, g->x_src, s);
    upload_vec(d_xtgt, g->x_tgt, s);
    upload_vec(d_xtype, g->x_type, s);
    upload_vec(d_xval, g->x_val, s);

    // ---- K2: nodes ----
    int code_bits = ct.n ? 2 * k : 0;
    code_bits = std::max(code_bits, bit_length(max_x));
    if (code_bits == 0) code_bits = 1;
    const bool bitmap_path = code_bits <= kBitmapMaxBits && !(flags & PDG_F_FORCE_SORT_NODES);
    NodeLookup L{nullptr, nullptr, nullptr, 0};
    if (bitmap_path) {
        const uint64_t n_words = std::max<uint64_t>(1, (1ULL << code_bits) >> 5);
        g->bitmap.reserve(n_words * 4, s);
        g->bitmap_prefix.reserve((n_words + 1) * 4, s);
        g->scratch0.reserve(n_words * 4, s);
        uint32_t *bm = g->bitmap.as<uint32_t>();
        PDG_CUDA(cudaMemsetAsync(bm, 0, n_words * 4, s));
        run_traverse<MODE_MARK>(g, ct, pos0, pos1, bm, nullptr, nullptr, L);
        if (!xcodes.empty()) {
            mark_codes_kernel<<<grid_for(xcodes.size(), 256), 256, 0, s>>>(d_xcodes.as<uint64_t>(), xcodes.size(), bm);
            PDG_LAUNCH_CHECK();
        }
        uint32_t *cnt = g->scratch0.as<uint32_t>();
        popc_words_kernel<<<grid_for(n_words, 256), 256, 0, s>>>(bm, n_words, cnt);
        PDG_LAUNCH_CHECK();
        exclusive_scan_u32_u32(cnt, g->bitmap_prefix.as<uint32_t>(), n_words, g);
        uint32_t nn = 0;
        PDG_CUDA(cudaMemcpyAsync(&nn, g->bitmap_prefix.as<uint32_t>() + n_words, 4, cudaMemcpyDeviceToHost, s));
        PDG_CUDA(cudaStreamSynchronize(s));
        g->n_nodes = nn;
        g->node_codes.reserve(std::max<size_t>((size_t)nn * 8, 16), s);
        bitmap_fill_codes_kernel<<<grid_for(n_words, 256), 256, 0, s>>>(bm, g->bitmap_prefix.as<uint32_t>(), n_words,
                                                                       g->node_codes.as<uint64_t>());
        PDG_LAUNCH_CHECK();
        g->have_bitmap = true;
        g->bitmap_words = n_words;
        L.bitmap = bm;
        L.prefix = g->bitmap_prefix.as<uint32_t>();
        st.node_path = 0;
    } else {
        const uint64_t n_codes = g->n_kmers + xcodes.size();
        g->have_bitmap = false;
        if (n_codes == 0) {
            g->n_nodes = 0;
            g->node_codes.reserve(16, s);
        } else {
            g->recA.reserve(n_codes * 8, s);
            g->recB.reserve(n_codes * 8, s);
            uint64_t *ka = g->recA.as<uint64_t>(), *kb = g->recB.as<uint64_t>();
            run_traverse<MODE_CODES>(g, ct, pos0, pos1, nullptr, ka, nullptr, L);
            if (!xcodes.empty())
                PDG_CUDA(cudaMemcpyAsync(ka + g->n_kmers, d_xcodes.p, xcodes.size() * 8, cudaMemcpyDeviceToDevice, s));
            uint64_t pr = 0;
            uint64_t *sorted = radix_sort_keys(g, ka, kb, n_codes, code_bits, &pr);
            g->scratch0.reserve(n_codes * 4, s);
            g->scratch1.reserve((n_codes + 1) * 8, s);
            uint32_t *flg = g->scratch0.as<uint32_t>();
            uint64_t *posn = g->scratch1.as<uint64_t>();
            head_flags_u64_kernel<<<grid_for(n_codes, 256), 256, 0, s>>>(sorted, n_codes, flg);
            PDG_LAUNCH_CHECK();
            exclusive_scan_u32_u64(flg, posn, n_codes, g);
            uint64_t nn = 0;
            PDG_CUDA(cudaMemcpyAsync(&nn, posn + n_codes, 8, cudaMemcpyDeviceToHost, s));
            PDG_CUDA(cudaStreamSynchronize(s));
            g->n_nodes = nn;
            g->node_codes.reserve(std::max<size_t>(nn * 8, 16), s);
            compact_u64_kernel<<<grid_for(n_codes, 256), 256, 0, s>>>(sorted, flg, posn, n_codes,
                                                                     g->node_codes.as<uint64_t>());
            PDG_LAUNCH_CHECK();
        }
        L.codes = g->node_codes.as<uint64_t>();
        st.node_path = 1;
    }
    if (g->world > 1) {
        // K5: hash-partitioned exchange of the node keys; afterwards every rank holds the same table
        if (!xcodes.empty()) {
            set_error("pdg_finalize: per-edge API content is not supported with world > 1");
            return PDG_E_INVALID;
        }
        const int rc = comm_global_nodes(g, 64);
        if (rc != PDG_OK) return rc;
        if (bitmap_path) {
            const uint64_t n_words = g->bitmap_words;
            uint32_t *bm = g->bitmap.as<uint32_t>();
            PDG_CUDA(cudaMemsetAsync(bm, 0, n_words * 4, s));
            if (g->n_nodes) {
                mark_codes_kernel<<<grid_for(g->n_nodes, 256), 256, 0, s>>>(g->node_codes.as<uint64_t>(), g->n_nodes, bm);
                PDG_LAUNCH_CHECK();
            }
            uint32_t *cnt = g->scratch0.as<uint32_t>();
            popc_words_kernel<<<grid_for(n_words, 256), 256, 0, s>>>(bm, n_words, cnt);
            PDG_LAUNCH_CHECK();
            exclusive_scan_u32_u32(cnt, g->bitmap_prefix.as<uint32_t>(), n_words, g);
        } else {
            L.codes = g->node_codes.as<uint64_t>();
        }
    }
    L.n_nodes = g->n_nodes;
    if (g->n_nodes >= (1ULL << 32)) {
        set_error("pdg_finalize: %llu nodes exceed 32-bit ids", (unsigned long long)g->n_nodes);
        return PDG_E_INVALID;
    }
    st.ms_nodes = lap.lap();

    // ---- K3: edge records ----
    const bool fast = g->contigs_sorted_fresh && n_x == 0 && !(flags & PDG_F_FORCE_GENERAL);
    st.general_sort = fast ? 0 : 1;
    g->row_ptr.reserve((g->n_nodes + 1) * 8, s);
    g->col_ptr.reserve((g->n_nodes + 1) * 8, s);
    g->n_edges = 0;
    if (n_rec == 0) {
        PDG_CUDA(cudaMemsetAsync(g->row_ptr.p, 0, (g->n_nodes + 1) * 8, s));
        PDG_CUDA(cudaMemsetAsync(g->col_ptr.p, 0, (g->n_nodes + 1) * 8, s));
    } else {
        g->recA.reserve(n_rec * sizeof(EdgeRec), s);
        g->recB.reserve(n_rec * sizeof(EdgeRec), s);
        EdgeRec *ra = g->recA.as<EdgeRec>(), *rb = g->recB.as<EdgeRec>();
        auto generate = [&]() {
            run_traverse<MODE_GEN>(g, ct, pos0, pos1, nullptr, nullptr, ra, L);
            if (n_x) {
                gen_explicit_kernel<<<grid_for(n_x, 256), 256, 0, s>>>(d_xsrc.as<uint64_t>(), d_xtgt.as<uint64_t>(),
                                                                      d_xtype.as<uint32_t>(), d_xval.as<int32_t>(), n_x,
                                                                      ra + g->n_gen_edges, L);
                PDG_LAUNCH_CHECK();
            }
        };
        const int id_bits = std::max(1, bit_length(g->n_nodes ? g->n_nodes - 1 : 0));
        uint32_t max_type = 0;
        for (uint32_t t : g->c_type) max_type = std::max(max_type, t);
        for (uint32_t t : g->x_type) max_type = std::max(max_type, t);
        const int type_bits = std::max(1, bit_length(max_type));
        auto key_passes = [&](int key_field) {
            std::vector<PassDesc> p;
            if (!fast) {   // least significant first: other endpoint, value, type
                auto a = field_passes(key_field == 0 ? 1 : 0, id_bits);
                auto b = field_passes(3, 32);
                auto c = field_passes(2, type_bits);
                p.insert(p.end(), a.begin(), a.end());
                p.insert(p.end(), b.begin(), b.end());
                p.insert(p.end(), c.begin(), c.end());
            }
            auto d = field_passes(key_field, id_bits);
            p.insert(p.end(), d.begin(), d.end());
            return p;
        };
        generate();
        st.ms_records = lap.lap();
        st.n_edge_records = n_rec;

        // CSR
        EdgeRec *sorted = radix_sort_records(g, ra, rb, n_rec, key_passes(0), &st.sort_passes_csr);
        uint64_t n_e = n_rec;
        if (!fast) {
            EdgeRec *other = sorted == ra ? rb : ra;
            g->scratch0.reserve(n_rec * 4, s);
            g->scratch1.reserve((n_rec + 1) * 8, s);
            uint32_t *flg = g->scratch0.as<uint32_t>();
            uint64_t *posn = g->scratch1.as<uint64_t>();
            head_flags_rec_kernel<<<grid_for(n_rec, 256), 256, 0, s>>>(sorted, n_rec, flg);
            PDG_LAUNCH_CHECK();
            exclusive_scan_u32_u64(flg, posn, n_rec, g);
            PDG_CUDA(cudaMemcpyAsync(&n_e, posn + n_rec, 8, cudaMemcpyDeviceToHost, s));
            PDG_CUDA(cudaStreamSynchronize(s));
            compact_rec_kernel<<<grid_for(n_rec, 256), 256, 0, s>>>(sorted, flg, posn, n_rec, other);
            PDG_LAUNCH_CHECK();
            sorted = other;
        }
        st.ms_sort_csr = lap.lap();
        g->n_edges = n_e;
        g->csr_tgt.reserve(n_e * 4, s);
        g->csr_type.reserve(n_e * 4, s);
        g->csr_val.reserve(n_e * 4, s);
        pack_rows_kernel<0><<<grid_for(n_e, 256), 256, 0, s>>>(sorted, n_e, g->n_nodes, g->row_ptr.as<uint64_t>(),
                                                              g->csr_tgt.as<uint32_t>(), g->csr_type.as<uint32_t>(),
                                                              g->csr_val.as<uint32_t>());
        PDG_LAUNCH_CHECK();
        st.ms_pack_csr = lap.lap();

        // CSC: fast path regenerates the records in (type,value) order; the general path re-sorts
        // the deduplicated CSR records on the full key
        EdgeRec *in2, *tmp2;
        if (fast) {
            generate();
            in2 = ra;
            tmp2 = rb;
        } else {
            in2 = sorted;
            tmp2 = sorted == ra ? rb : ra;
        }
        EdgeRec *sorted2 = radix_sort_records(g, in2, tmp2, n_e, key_passes(1), &st.sort_passes_csc);
        st.ms_sort_csc = lap.lap();
        g->csc_src.reserve(n_e * 4, s);
        g->csc_type.reserve(n_e * 4, s);
        g->csc_val.reserve(n_e * 4, s);
        pack_rows_kernel<1><<<grid_for(n_e, 256), 256, 0, s>>>(sorted2, n_e, g->n_nodes, g->col_ptr.as<uint64_t>(),
                                                              g->csc_src.as<uint32_t>(), g->csc_type.as<uint32_t>(),
                                                              g->csc_val.as<uint32_t>());
        PDG_LAUNCH_CHECK();
        st.ms_pack_csc = lap.lap();
    }
    PDG_CUDA(cudaStreamSynchronize(s));
    st.ms_total = total.lap();
    st.n_residues = g->n_res;
    st.n_contigs = g->c_begin.size();
    st.n_kmers = g->n_kmers;
    st.n_nodes = g->n_nodes;
    st.n_edges = g->n_edges;
    st.ms_onesweep_avg = g->stats.ms_onesweep_avg;
    st.onesweep_launches = g->stats.onesweep_launches;
    st.onesweep_records = g->stats.onesweep_records;
    g->stats = st;
    g->finalized = true;
    return PDG_OK;
}

int kmer_stats(pdg_graph *g, const uint64_t *h_begin, const uint64_t *h_len, uint64_t n, uint64_t *n_kmers,
               uint64_t *n_unique) {
    cudaStream_t s = g->stream;
    const int k = g->k;
    std::vector<uint64_t> begin, len, ebase;
    std::vector<uint32_t> type;
    uint64_t e = 0, km = 0;
    for (uint64_t i = 0; i < n; ++i) {
        if (h_len[i] <= (uint64_t)k) continue;
        if (h_begin[i] + h_len[i] > g->n_res) {
            set_error("pdg_kmer_stats: contig beyond the residue store");
            return PDG_E_INVALID;
        }
        begin.push_back(h_begin[i]);
        len.push_back(h_len[i]);
        ebase.push_back(e);
        type.push_back(0);
        e += h_len[i] - k;
        km += h_len[i] - k + 1;
    }
    *n_kmers = km;
    *n_unique = 0;
    if (km == 0) return PDG_OK;
    for (size_t i = 1; i < begin.size(); ++i)
        if (begin[i] < begin[i - 1] + len[i - 1]) {
            set_error("pdg_kmer_stats: contigs must be ascending and disjoint");
            return PDG_E_INVALID;
        }
    DevBuf db, de, dx, dt, dsum;
    ContigTable ct = make_table(begin, len, ebase, type, db, de, dx, dt, s);
    uint64_t pos0, pos1;
    table_span(begin, len, &pos0, &pos1);
    dsum.reserve(16, s);
    PDG_CUDA(cudaMemsetAsync(dsum.p, 0, 16, s));
    NodeLookup L{nullptr, nullptr, nullptr, 0};
    if (2 * k <= kBitmapMaxBits) {
        const uint64_t n_words = std::max<uint64_t>(1, (1ULL << (2 * k)) >> 5);
        DevBuf bm;
        bm.reserve(n_words * 4, s);
        PDG_CUDA(cudaMemsetAsync(bm.p, 0, n_words * 4, s));
        run_traverse<MODE_MARK>(g, ct, pos0, pos1, bm.as<uint32_t>(), nullptr, nullptr, L);
        count_set_bits_kernel<<<grid_for(n_words, 256), 256, 0, s>>>(bm.as<uint32_t>(), n_words,
                                                                    dsum.as<unsigned long long>());
        PDG_LAUNCH_CHECK();
        PDG_CUDA(cudaMemcpyAsync(n_unique, dsum.p, 8, cudaMemcpyDeviceToHost, s));
        PDG_CUDA(cudaStreamSynchronize(s));
    } else {
        DevBuf ka, kb;
        ka.reserve(km * 8, s);
        kb.reserve(km * 8, s);
        run_traverse<MODE_CODES>(g, ct, pos0, pos1, nullptr, ka.as<uint64_t>(), nullptr, L);
        uint64_t pr = 0;
        uint64_t *sorted = radix_sort_keys(g, ka.as<uint64_t>(), kb.as<uint64_t>(), km, 2 * k, &pr);
        count_heads_kernel<<<grid_for(km, 256), 256, 0, s>>>(sorted, km, dsum.as<unsigned long long>());
        PDG_LAUNCH_CHECK();
        PDG_CUDA(cudaMemcpyAsync(n_unique, dsum.p, 8, cudaMemcpyDeviceToHost, s));
        PDG_CUDA(cudaStreamSynchronize(s));
    }
    return PDG_OK;
}

int emit_codes(pdg_graph *g, uint64_t begin, uint64_t len, uint64_t *d_out) {
    cudaStream_t s = g->stream;
    if (len < (uint64_t)g->k) return PDG_OK;
    if (begin + len > g->n_res) {
        set_error("pdg_emit_codes: range beyond the residue store");
        return PDG_E_INVALID;
    }
    std::vector<uint64_t> b{begin}, l{len}, e{0};
    std::vector<uint32_t> t{0};
    DevBuf db, de, dx, dt;
    ContigTable ct = make_table(b, l, e, t, db, de, dx, dt, s);
    NodeLookup L{nullptr, nullptr, nullptr, 0};
    run_traverse<MODE_CODES>(g, ct, begin, begin + len, nullptr, d_out, nullptr, L);
    PDG_CUDA(cudaStreamSynchronize(s));
    return PDG_OK;
}

int checksum(pdg_graph *g, uint64_t *nodes_sum, uint64_t *edges_sum) {
    cudaStream_t s = g->stream;
    DevBuf d;
    d.reserve(32, s);
    PDG_CUDA(cudaMemsetAsync(d.p, 0, 32, s));
    unsigned long long *sums = d.as<unsigned long long>();
    if (g->n_nodes) {
        checksum_nodes_kernel<<<grid_for(g->n_nodes, 256), 256, 0, s>>>(g->node_codes.as<uint64_t>(), g->n_nodes, sums);
        PDG_LAUNCH_CHECK();
    }
    if (g->n_edges) {
        checksum_edges_kernel<<<grid_for(g->n_edges, 256), 256, 0, s>>>(
            g->row_ptr.as<uint64_t>(), g->n_nodes, g->n_edges, g->csr_tgt.as<uint32_t>(), g->csr_type.as<uint32_t>(),
            g->csr_val.as<uint32_t>(), g->node_codes.as<uint64_t>(), 1, sums + 1);
        PDG_LAUNCH_CHECK();
        checksum_edges_kernel<<<grid_for(g->n_edges, 256), 256, 0, s>>>(
            g->col_ptr.as<uint64_t>(), g->n_nodes, g->n_edges, g->csc_src.as<uint32_t>(), g->csc_type.as<uint32_t>(),
            g->csc_val.as<uint32_t>(), g->node_codes.as<uint64_t>(), 0, sums + 2);
        PDG_LAUNCH_CHECK();
    }
    unsigned long long h[3];
    PDG_CUDA(cudaMemcpyAsync(h, sums, 24, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    if (h[1] != h[2]) {
        set_error("pdg_checksum: CSR and CSC disagree (%llx vs %llx)", h[1], h[2]);
        return PDG_E_INVALID;
    }
    *nodes_sum = h[0];
    *edges_sum = h[1];
    return PDG_OK;
}
}  // namespace pdg
