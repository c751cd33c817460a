// gbuild.cu -- the k-mer occurrence index of the genome part.
//   records (code, position) in emission order  --stable LSD radix sort on code-->  grouped by
//   code with ascending positions  --run heads-->  rows (= distinct k-mers = nodes, dense id =
//   rank) + row pointers + the position column.
// Reference counterpart: the node_idx / srcnode_idx / tgtnode_idx B-trees that LemonGraph fills
// one key at a time (lemongraph/lib/lemongraph.c:591-630).
#include "gindex.cuh"
#include "msd3.cuh"
#include <algorithm>

namespace pdg {

template <typename Rec> __device__ __forceinline__ uint64_t rec_code(const Rec &r);
template <> __device__ __forceinline__ uint64_t rec_code<uint64_t>(const uint64_t &r) { return r >> 32; }
template <> __device__ __forceinline__ uint64_t rec_code<uint4>(const uint4 &r) { return ((uint64_t)r.y << 32) | r.x; }
template <typename Rec> __device__ __forceinline__ uint32_t rec_pos(const Rec &r);
template <> __device__ __forceinline__ uint32_t rec_pos<uint64_t>(const uint64_t &r) { return (uint32_t)r; }
template <> __device__ __forceinline__ uint32_t rec_pos<uint4>(const uint4 &r) { return r.z; }

template <typename Rec>
__global__ void g_head_flags_kernel(const Rec *__restrict__ r, uint64_t n, uint32_t *__restrict__ flags) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = (i == 0 || rec_code(r[i]) != rec_code(r[i - 1])) ? 1u : 0u;
}

template <typename Rec>
__global__ void g_rows_kernel(const Rec *__restrict__ r, const uint32_t *__restrict__ flags,
                              const uint32_t *__restrict__ rank, uint64_t n, uint64_t *__restrict__ codes,
                              uint32_t *__restrict__ rowptr, uint32_t *__restrict__ pos) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Rec x = r[i];
    pos[i] = rec_pos(x);
    if (flags[i]) {
        const uint32_t id = rank[i];
        codes[id] = rec_code(x);
        rowptr[id] = (uint32_t)i;
    }
    if (i == n - 1) rowptr[rank[n]] = (uint32_t)n;
}

__global__ void g_mark_codes_kernel(const uint64_t *__restrict__ codes, uint64_t n, uint32_t *__restrict__ bitmap) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t c = codes[i];
    atomicOr(&bitmap[c >> 5], 1u << ((uint32_t)c & 31u));
}

__global__ void g_popc_words_kernel(const uint32_t *__restrict__ bm, uint64_t n_words, uint32_t *__restrict__ cnt) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_words) cnt[i] = __popc(bm[i]);
}

// head_flags: one word per record, 1 where a new code starts (computed here when null)
template <typename Rec>
static void rows_from_sorted(pdg_graph *g, const Rec *sorted, uint64_t n, const uint32_t *head_flags = nullptr) {
    cudaStream_t s = g->stream;
    g->scratch1.reserve((n + 1) * 4, s);
    uint32_t *rank = g->scratch1.as<uint32_t>();
    const uint32_t *flg = head_flags;
    if (!flg) {
        g->scratch0.reserve(n * 4, s);
        g_head_flags_kernel<Rec><<<grid_for(n, 256), 256, 0, s>>>(sorted, n, g->scratch0.as<uint32_t>());
        PDG_LAUNCHED(g);
        flg = g->scratch0.as<uint32_t>();
    }
    exclusive_scan_u32_u32(flg, rank, n, g);
    uint32_t nn = 0;
    PDG_CUDA(cudaMemcpyAsync(&nn, rank + n, 4, cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    g->g_nodes = nn;
    g->g_codes.reserve(std::max<size_t>((size_t)nn * 8, 16), s);
    g->g_rowptr.reserve(((size_t)nn + 1) * 4, s);
    g->g_pos.reserve(n * 4, s);
    g_rows_kernel<Rec><<<grid_for(n, 256), 256, 0, s>>>(sorted, flg, rank, n, g->g_codes.as<uint64_t>(),
                                                       g->g_rowptr.as<uint32_t>(), g->g_pos.as<uint32_t>());
    PDG_LAUNCHED(g);
}

void build_g_index(pdg_graph *g, const ContigTable &ct, const HostContigs &hc, int flags) {
    cudaStream_t s = g->stream;
    const int k = g->k;
    const uint64_t n = hc.n_kmers;
    g->g_kmers = n;
    g->g_edges = n - hc.size();
    g->g_nodes = 0;
    g->have_bitmap = false;
    g->g_rowptr.reserve(16, s);
    PDG_CUDA(cudaMemsetAsync(g->g_rowptr.p, 0, 16, s));
    g->g_codes.reserve(16, s);
    if (n == 0) return;
    TraverseOut out;
    uint64_t passes = 0;
    g->collect_sort_stats = true;
    // a pipelined ingest already ran passes 1-2 (and launched pass 3) group by group -- if the caller
    // declared exactly the contigs it assumed, only the rest is left
    bool staged = g->pre_valid && !g->m3_groups.empty() && g->pre_begin.size() == hc.size() &&
                  !(flags & PDG_F_FORCE_SORT_NODES);
    for (size_t i = 0; staged && i < hc.size(); ++i)
        staged = g->pre_begin[i] == hc.begin[i] && g->pre_begin[i] + g->pre_len[i] == hc.end[i];
    g->stats.pipelined = staged ? 1 : 0;
    if (k <= 16) {
        // fused path (ksort.cu): the first pass reads the packed genome, the last writes positions
        const uint64_t *sorted = nullptr;
        if (2 * k <= kMsd3MaxBits && !(flags & PDG_F_FORCE_SORT_NODES)) {   // msd3.cu
            if (staged) msd3_finish(g, n); else msd3_build(g, ct, hc);
        }
        else if (!ksort_build(g, ct, hc, &sorted)) rows_from_sorted<uint64_t>(g, sorted, n);
        passes = g->stats.sort_passes;
    } else if (!(flags & PDG_F_FORCE_SORT_NODES) && msdlong_build(g, ct, hc, staged)) {   // msdlong.cu: prefix MSD + bucket sort + rows
        passes = g->stats.sort_passes;
    } else {
        g->recA.reserve(n * 16, s);
        g->recB.reserve(n * 16, s);
        out.rec = g->recA.as<uint4>();
        run_traverse(g, MODE_REC128, ct, hc.pos0, hc.pos1, out);
        std::vector<PassDesc> p;
        for (int sft = 0; sft < 32; sft += 8) p.push_back(PassDesc{0, sft});
        for (int sft = 0; sft < 2 * k - 32; sft += 8) p.push_back(PassDesc{1, sft});
        const uint4 *sorted = radix_sort_records(g, g->recA.as<uint4>(), g->recB.as<uint4>(), n, p, &passes);
        rows_from_sorted<uint4>(g, sorted, n);
    }
    g->collect_sort_stats = false;
    g->pre_valid = false;
    g->m3_groups.clear();
    g->m3_launched = false;
    g->stats.sort_passes = passes;
    g->stats.sort_records = n;
    // O(1) lookup for small k: presence bitmap + popcount prefix (rank == row id)
    if (2 * k <= kBitmapMaxBits && !(flags & PDG_F_FORCE_SORT_NODES)) {
        const uint64_t n_words = std::max<uint64_t>(1, (1ULL << (2 * k)) >> 5);
        g->bitmap.reserve(n_words * 4, s);
        g->bitmap_prefix.reserve((n_words + 1) * 4, s);
        g->scratch0.reserve(n_words * 4, s);
        uint32_t *bm = g->bitmap.as<uint32_t>();
        PDG_CUDA(cudaMemsetAsync(bm, 0, n_words * 4, s));
        g_mark_codes_kernel<<<grid_for(g->g_nodes, 256), 256, 0, s>>>(g->g_codes.as<uint64_t>(), g->g_nodes, bm);
        PDG_LAUNCHED(g);
        g_popc_words_kernel<<<grid_for(n_words, 256), 256, 0, s>>>(bm, n_words, g->scratch0.as<uint32_t>());
        PDG_LAUNCHED(g);
        exclusive_scan_u32_u32(g->scratch0.as<uint32_t>(), g->bitmap_prefix.as<uint32_t>(), n_words, g);
        g->have_bitmap = true;
        g->bitmap_words = n_words;
    }
}

ContigTable g_table(const pdg_graph *g) {
    ContigTable ct;
    ct.begin = g->d_gbegin.as<uint64_t>();
    ct.end = g->d_gend.as<uint64_t>();
    ct.kbase = g->d_gkbase.as<uint64_t>();
    ct.type = g->d_gtype.as<uint32_t>();
    ct.n = g->g_begin.size();
    return ct;
}

NodeLookup g_lookup(const pdg_graph *g) {
    NodeLookup L;
    L.bitmap = g->have_bitmap ? g->bitmap.as<uint32_t>() : nullptr;
    L.prefix = g->bitmap_prefix.as<uint32_t>();
    L.codes = g->g_codes.as<uint64_t>();
    L.n_nodes = g->g_nodes;
    L.bitmap_bits = g->bitmap_words * 32;
    return L;
}

}  // namespace pdg
