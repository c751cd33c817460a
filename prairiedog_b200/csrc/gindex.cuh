// gindex.cuh -- host-side declarations of the genome-part build (gindex.cu, gbuild.cu).
#pragma once
#include "traverse.cuh"

namespace pdg {

static constexpr int EM_THREADS = 256;
static constexpr int EM_ITEMS = 8;
static constexpr int EM_TILE = EM_THREADS * EM_ITEMS;      // residue positions per CTA
static constexpr int EM_WORDS = EM_TILE / 32 + 2;          // packed words staged per CTA

enum { MODE_MARK = 0, MODE_CODES = 1, MODE_REC64 = 2, MODE_REC128 = 3, MODE_XGEN = 4 };

struct TraverseOut {
    uint32_t *bitmap = nullptr;     // MARK
    uint64_t *codes = nullptr;      // CODES: code; REC64: (code << 32) | position
    uint4 *rec = nullptr;           // REC128: {code lo, code hi, position, 0}
    uint64_t *x_src = nullptr, *x_tgt = nullptr;   // XGEN
    uint32_t *x_type = nullptr;
    int32_t *x_val = nullptr;
};

// host image of a contig table (every contig has len > k, ascending, disjoint)
struct HostContigs {
    std::vector<uint64_t> begin, end, kbase;
    std::vector<uint32_t> type;
    uint64_t n_kmers = 0, pos0 = 0, pos1 = 0;
    HostContigs() = default;
    HostContigs(const std::vector<uint64_t> &b, const std::vector<uint64_t> &l, const std::vector<uint32_t> &t, int k);
    ContigTable upload(DevBuf &db, DevBuf &de, DevBuf &dk, DevBuf &dt, cudaStream_t s) const;
    size_t size() const { return begin.size(); }
};

void run_traverse(pdg_graph *g, int mode, const ContigTable &ct, uint64_t pos0, uint64_t pos1,
                  const TraverseOut &out);

// gbuild.cu: builds g->g_codes / g_rowptr / g_pos (+ bitmap lookup) from the contigs in g->g_*
void build_g_index(pdg_graph *g, const ContigTable &ct, const HostContigs &hc, int flags);
ContigTable g_table(const pdg_graph *g);
NodeLookup g_lookup(const pdg_graph *g);

// ksort.cu: fused emission + sort for k <= 16; true when rows were built from the direct table
bool ksort_build(pdg_graph *g, const ContigTable &ct, const HostContigs &hc, const uint64_t **sorted_out);

// msd3.cu: offsets-first MSD sort for k <= 11 (no look-back)
void msd3_build(pdg_graph *g, const ContigTable &ct, const HostContigs &hc);
// the same build in steps (groups of positions staged one by one, ingest_batch.cu)
void msd3_reserve(pdg_graph *g, uint64_t pos_lo, uint64_t pos_hi, uint32_t max_groups);
void msd3_stage_group(pdg_graph *g, const ContigTable &ct, uint64_t pos0, uint64_t pos1);
void msd3_finish_launch(pdg_graph *g, bool rows = true);                 // pass 3 + presence + ranks, asynchronous
void msd3_finish_complete(pdg_graph *g, uint64_t n_kmers);
void msd3_finish(pdg_graph *g, uint64_t n_kmers);      // launch (unless already launched) + complete

// msdlong.cu: 17 <= k <= 31: msd3 on the 11-residue prefix + a sort of every bucket in registers; on success the
// occurrence index (g_pos, g_codes, g_rowptr, g_nodes) is complete;
// false = not applicable / a bucket did not fit (the caller runs the LSD sort of radix.cu)
bool msdlong_build(pdg_graph *g, const ContigTable &ct, const HostContigs &hc, bool staged);

// xbuild.cu: builds the explicit part from g->x_* plus the demoted contigs
void build_x_part(pdg_graph *g, const HostContigs &demoted);

static inline unsigned grid_for(uint64_t n, int threads) { return (unsigned)((n + threads - 1) / threads); }
static inline int bit_length(uint64_t v) { int b = 0; while (v) { ++b; v >>= 1; } return b; }

}  // namespace pdg
