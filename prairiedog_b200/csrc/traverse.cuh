// traverse.cuh -- device helpers shared by the build, export and query kernels: the contig table,
// k-mer extraction from the 2-bit store, node lookup.
//   prairiedog/kmers.py:96-124         sliding window            -> kmer_at / kmer_global
//   prairiedog/subgraph_ref.py:62-102  edge (kmer_i -> kmer_i+1) -> position p of contig c
#pragma once
#include "common.cuh"

namespace pdg {

struct ContigTable {
    const uint64_t *begin;   // ascending
    const uint64_t *end;
    const uint64_t *kbase;   // k-mers before this contig (exclusive prefix of len-k+1)
    const uint32_t *type;
    uint64_t n;
};

// rows of the occurrence index: sorted unique codes, optionally a presence bitmap for O(1) rank
struct NodeLookup {
    const uint32_t *bitmap;      // presence bits or nullptr
    const uint32_t *prefix;      // exclusive popcount prefix per 32-bit word
    const uint64_t *codes;       // sorted unique codes
    uint64_t n_nodes;
    uint64_t bitmap_bits;        // codes >= this are never in the bitmap
};

#ifdef __CUDACC__
// index of `code` in L.codes, or -1
__device__ __forceinline__ int64_t find_code(const NodeLookup &L, uint64_t code) {
    if (L.bitmap) {
        if (code >= L.bitmap_bits) return -1;
        const uint32_t w = (uint32_t)(code >> 5), b = (uint32_t)code & 31u;
        const uint32_t m = L.bitmap[w];
        if (!((m >> b) & 1u)) return -1;
        return (int64_t)(L.prefix[w] + __popc(m & ((1u << b) - 1u)));
    }
    uint64_t lo = 0, hi = L.n_nodes;   // lower_bound
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (L.codes[mid] < code) lo = mid + 1; else hi = mid;
    }
    return (lo < L.n_nodes && L.codes[lo] == code) ? (int64_t)lo : -1;
}

// last contig whose begin <= p (0 when p precedes every contig)
__device__ __forceinline__ uint64_t contig_of(const ContigTable &ct, uint64_t p) {
    uint64_t lo = 0, hi = ct.n;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (ct.begin[mid] <= p) lo = mid + 1; else hi = mid;
    }
    return lo ? lo - 1 : 0;
}

// k-mer starting at residue slot q of a word array (big endian 2-bit packing, 32 per word);
// reads words[q>>5] and words[(q>>5)+1]
__device__ __forceinline__ uint64_t kmer_at(const unsigned long long *words, uint64_t q, int k) {
    const uint64_t w = q >> 5;
    const uint32_t s = 2u * (uint32_t)(q & 31u);
    const unsigned long long hi = words[w], lo = words[w + 1];
    const unsigned long long x = s ? ((hi << s) | (lo >> (64 - s))) : hi;
    return x >> (64 - 2 * k);
}

__device__ __forceinline__ uint32_t base_at(const unsigned long long *words, uint64_t q) {
    return (uint32_t)(words[q >> 5] >> (62 - 2 * (uint32_t)(q & 31u))) & 3u;
}
#endif

}  // namespace pdg
