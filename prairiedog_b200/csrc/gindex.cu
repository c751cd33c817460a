// gindex.cu -- K1b/K2/K3 for the genome part: k-mer emission from the packed store and the k-mer
// occurrence index (positions grouped by code, ascending inside a group).
//
// Restates, as bulk data-parallel passes, what the reference does one edge at a time:
//   prairiedog/kmers.py:96-124            sliding window  -> traverse_kernel
//   lemongraph/lib/lemongraph.c:661-672   node upsert unique on (type,value) -> run heads of the
//                                         sorted (code, position) records = the rows
//   lemongraph/lib/lemongraph.c:603-630   srcnode_idx / tgtnode_idx -> ONE index: an occurrence
//                                         at p is an out-edge unless it is the last k-mer of its
//                                         contig and an in-edge unless it is the first
#include "traverse.cuh"
#include "gindex.cuh"
#include <algorithm>

namespace pdg {

// One pass over residue positions [pos0, pos1).  Position p inside contig c (begin <= p,
// p + k <= end) is k-mer number p - begin of that contig.
template <int MODE>
__global__ void __launch_bounds__(EM_THREADS) traverse_kernel(const unsigned long long *__restrict__ packed,
                                                              uint64_t packed_words, uint64_t pos0,
                                                              uint64_t pos1, int k, ContigTable ct,
                                                              TraverseOut out) {
    __shared__ unsigned long long words[EM_WORDS];
    __shared__ uint64_t ci0_s;
    const uint64_t tile_p0 = pos0 + (uint64_t)blockIdx.x * EM_TILE;
    const uint64_t w0 = tile_p0 >> 5;
    for (int i = threadIdx.x; i < EM_WORDS; i += EM_THREADS)
        words[i] = (w0 + i < packed_words) ? packed[w0 + i] : 0ULL;
    if (threadIdx.x == 0) ci0_s = contig_of(ct, tile_p0);
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
        const uint64_t ord = ct.kbase[ci] + (p - cb);   // k-mer ordinal
        if (MODE == MODE_MARK) {
            const uint32_t w = (uint32_t)(code >> 5), bit = 1u << ((uint32_t)code & 31u);
            if (!(out.bitmap[w] & bit)) atomicOr(&out.bitmap[w], bit);
        } else if (MODE == MODE_CODES) {
            out.codes[ord] = code;
        } else if (MODE == MODE_REC64) {
            out.codes[ord] = (code << 32) | p;
        } else if (MODE == MODE_REC128) {
            out.rec[ord] = make_uint4((uint32_t)code, (uint32_t)(code >> 32), (uint32_t)p, 0u);
        } else {   // MODE_XGEN: explicit edge records of a demoted contig; edge ordinal = ord - ci
            if (p + k < ce) {
                const uint64_t mask = (1ULL << (2 * k)) - 1ULL;
                const uint64_t tcode = ((code << 2) | base_at(words, lead + l + k)) & mask;
                const uint64_t e = ord - ci;
                out.x_src[e] = code;
                out.x_tgt[e] = tcode;
                out.x_type[e] = ct.type[ci];
                out.x_val[e] = (int32_t)(p - cb);
            }
        }
    }
}

template <int MODE>
static void launch_traverse(pdg_graph *g, const ContigTable &ct, uint64_t pos0, uint64_t pos1,
                            const TraverseOut &out) {
    if (ct.n == 0 || pos1 <= pos0) return;
    const uint64_t n_tiles = (pos1 - pos0 + EM_TILE - 1) / EM_TILE;
    traverse_kernel<MODE><<<(unsigned)n_tiles, EM_THREADS, 0, g->stream>>>(
        g->packed.as<unsigned long long>(), g->packed.cap / 8, pos0, pos1, g->k, ct, out);
    PDG_LAUNCHED(g);
}

void run_traverse(pdg_graph *g, int mode, const ContigTable &ct, uint64_t pos0, uint64_t pos1,
                  const TraverseOut &out) {
    switch (mode) {
        case MODE_MARK: launch_traverse<MODE_MARK>(g, ct, pos0, pos1, out); break;
        case MODE_CODES: launch_traverse<MODE_CODES>(g, ct, pos0, pos1, out); break;
        case MODE_REC64: launch_traverse<MODE_REC64>(g, ct, pos0, pos1, out); break;
        case MODE_REC128: launch_traverse<MODE_REC128>(g, ct, pos0, pos1, out); break;
        default: launch_traverse<MODE_XGEN>(g, ct, pos0, pos1, out); break;
    }
}

HostContigs::HostContigs(const std::vector<uint64_t> &b, const std::vector<uint64_t> &l,
                         const std::vector<uint32_t> &t, int k)
    : begin(b), type(t) {
    end.resize(b.size());
    kbase.resize(b.size());
    uint64_t kb = 0;
    pos0 = b.empty() ? 0 : b.front();
    pos1 = 0;
    for (size_t i = 0; i < b.size(); ++i) {
        end[i] = b[i] + l[i];
        kbase[i] = kb;
        kb += l[i] - (uint64_t)k + 1;
        pos1 = std::max(pos1, end[i]);
    }
    n_kmers = kb;
}

template <typename T>
static void upload_vec(DevBuf &buf, const std::vector<T> &v, cudaStream_t s) {
    buf.reserve(std::max<size_t>(v.size() * sizeof(T), 16), s);
    if (!v.empty()) PDG_CUDA(cudaMemcpyAsync(buf.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, s));
}

ContigTable HostContigs::upload(DevBuf &db, DevBuf &de, DevBuf &dk, DevBuf &dt, cudaStream_t s) const {
    upload_vec(db, begin, s);
    upload_vec(de, end, s);
    upload_vec(dk, kbase, s);
    upload_vec(dt, type, s);
    PDG_CUDA(cudaStreamSynchronize(s));   // pageable sources
    ContigTable ct;
    ct.begin = db.as<uint64_t>();
    ct.end = de.as<uint64_t>();
    ct.kbase = dk.as<uint64_t>();
    ct.type = dt.as<uint32_t>();
    ct.n = begin.size();
    return ct;
}

}  // namespace pdg
