// ingest.cu -- K0/K1a: FASTA bytes on the device -> 2-bit packed residues + header table.
//
// Restates Kmers._load (prairiedog/kmers.py:51-77) as a data-parallel scan.  A file is a string
// over a three state line automaton (LS = at line start, H = inside a '>' line, S = inside a
// sequence line).  Every thread summarises 64 bytes as a map {entry state -> exit state, residue
// count, header count, dirty count}; maps compose associatively, so a block scan plus a scan over
// the 16 KiB tiles gives every thread its true entry state and its residue / header offsets.
//   pass A  fasta_tile_summaries : 1 B/byte read  -> one summary per tile
//   pass B  fasta_tile_scan      : single CTA over the tile summaries
//   pass C  fasta_emit           : 1 B/byte read  -> 0.25 B/residue packed store, header table
// rstrip (kmers.py:63) is honoured for '\r', ' ', '\t', '\v', '\f' at line ends.  Anything the
// 2-bit alphabet cannot hold (bytes outside ACGT in a sequence line, blanks inside a line, a CR
// that is not followed by LF) is counted in n_dirty and the caller routes that file through the
// per-edge path instead, so the result stays exact.
#include "common.cuh"

namespace pdg {

static constexpr int IN_THREADS = 256;
static constexpr int IN_CHUNK = 64;                       // bytes per thread
static constexpr int IN_TILE = IN_THREADS * IN_CHUNK;     // 16 KiB per CTA
static constexpr int IN_ROW = IN_CHUNK / 4 + 1;           // padded row (words): conflict free

enum { LS = 0, HD = 1, SQ = 2 };

struct Summ {
    uint32_t res[3], hdr[3], dirty[3];
    uint32_t out;   // bits [2e, 2e+2) = exit state for entry state e
};

__device__ __forceinline__ Summ summ_identity() {
    Summ s;
#pragma unroll
    for (int e = 0; e < 3; ++e) s.res[e] = s.hdr[e] = s.dirty[e] = 0;
    s.out = (LS) | (HD << 2) | (SQ << 4);
    return s;
}

// first A then B
__device__ __forceinline__ Summ summ_compose(const Summ &A, const Summ &B) {
    Summ c;
    c.out = 0;
#pragma unroll
    for (int e = 0; e < 3; ++e) {
        const uint32_t s = (A.out >> (2 * e)) & 3u;
        c.res[e] = A.res[e] + B.res[s];
        c.hdr[e] = A.hdr[e] + B.hdr[s];
        c.dirty[e] = A.dirty[e] + B.dirty[s];
        c.out |= ((B.out >> (2 * s)) & 3u) << (2 * e);
    }
    return c;
}

__device__ __forceinline__ bool is_blank(int c) {
    return c == ' ' || c == '\t' || c == '\r' || c == 0x0b || c == 0x0c;
}
__device__ __forceinline__ bool is_acgt(int c) { return c == 'A' || c == 'C' || c == 'G' || c == 'T'; }
// A=0 T=1 C=2 G=3 : the order of itertools.product('ATCG') in kmers.py:23
__device__ __forceinline__ uint32_t base_code(int c) {
    const uint32_t x = ((uint32_t)c >> 1) & 3u;   // A0 C1 T2 G3
    return ((x & 1u) << 1) | (x >> 1);
}

struct WalkOut { uint32_t res, hdr, dirty, state; };

// Emission context of pass C (unused by pass A).
struct EmitCtx {
    unsigned long long *words;   // shared packed words of this tile
    uint32_t q;                  // residue slot inside the tile buffer (includes base & 31)
    uint64_t res_global;         // residues before this thread (global, file relative)
    uint64_t hdr_global;         // headers started before this thread
    uint64_t byte_base;          // file offset of this thread's first byte
    uint64_t *table;             // rows {begin, end, residue_index}
    uint64_t table_cap;
};

// Runs the line automaton over one thread's bytes (in shared memory, padded rows).
// next_c: the byte after the chunk, -1 at end of file.
template <bool EMIT>
__device__ __forceinline__ WalkOut walk_chunk(const uint32_t *row, int len, int next_c, uint32_t state,
                                              EmitCtx *ctx) {
    WalkOut o{0u, 0u, 0u, state};
    unsigned long long acc = 0;
    int acc_word = -1;
    int c = len > 0 ? (int)(row[0] & 0xffu) : -1;
    for (int i = 0; i < len; ++i) {
        const int j = i + 1;
        const int nx = j < len ? (int)((row[j >> 2] >> ((j & 3) * 8)) & 0xffu) : next_c;
        uint32_t st = o.state;
        if (st == LS) {
            if (c == '>') {
                st = HD;
                if (EMIT) {
                    const uint64_t h = ctx->hdr_global + o.hdr;
                    if (h < ctx->table_cap) {
                        ctx->table[3 * h + 0] = ctx->byte_base + i;
                        ctx->table[3 * h + 2] = ctx->res_global + o.res;
                    }
                }
                o.hdr += 1;
                o.state = st;
                c = nx;
                continue;
            }
            st = SQ;   // falls through: the byte is handled as a sequence-line byte
        }
        if (c == '\n') {
            if (EMIT && st == HD) {
                const uint64_t h = ctx->hdr_global + o.hdr - 1;   // header this line belongs to
                if (h < ctx->table_cap) ctx->table[3 * h + 1] = ctx->byte_base + i;
            }
            st = LS;
        } else if (st == HD) {
            if (c == '\r' && !(nx == '\n' || nx < 0)) o.dirty += 1;
        } else {   // sequence line
            if (is_acgt(c)) {
                if (EMIT) {
                    const uint32_t q = ctx->q + o.res;
                    const int wd = (int)(q >> 5);
                    if (wd != acc_word) {
                        if (acc_word >= 0) atomicOr(&ctx->words[acc_word], acc);
                        acc = 0;
                        acc_word = wd;
                    }
                    acc |= (unsigned long long)base_code(c) << (62 - 2 * (q & 31u));
                }
                o.res += 1;
            } else if (c == '\r') {
                if (!(nx == '\n' || nx < 0)) { o.dirty += 1; o.res += 1; }
            } else if (is_blank(c)) {
                if (!(nx == '\n' || nx < 0 || is_blank(nx))) { o.dirty += 1; o.res += 1; }
            } else {
                o.dirty += 1;
                o.res += 1;
            }
        }
        o.state = st;
        c = nx;
    }
    if (EMIT && acc_word >= 0) atomicOr(&ctx->words[acc_word], acc);
    return o;
}

// Loads the tile into padded shared rows with 128-bit coalesced loads; returns bytes in tile.
__device__ __forceinline__ uint32_t load_tile(const uint8_t *__restrict__ bytes, uint64_t n, uint64_t tile_base,
                                              uint32_t *rows) {
    const uint32_t tile_n = (uint32_t)min((uint64_t)IN_TILE, n - tile_base);
    const bool aligned = ((reinterpret_cast<uintptr_t>(bytes) + tile_base) & 15) == 0;
    for (uint32_t v = threadIdx.x; v < IN_TILE / 16; v += IN_THREADS) {
        const uint32_t off = v * 16;
        uint4 x = make_uint4(0, 0, 0, 0);
        if (aligned && off + 16 <= tile_n) {
            x = __ldg(reinterpret_cast<const uint4 *>(bytes + tile_base + off));
        } else if (off < tile_n) {
            uint32_t wv[4] = {0, 0, 0, 0};
            for (uint32_t b = 0; b < 16 && off + b < tile_n; ++b)
                wv[b >> 2] |= (uint32_t)bytes[tile_base + off + b] << ((b & 3) * 8);
            x = make_uint4(wv[0], wv[1], wv[2], wv[3]);
        }
        const uint32_t t = off / IN_CHUNK, wcol = (off % IN_CHUNK) / 4;
        uint32_t *dst = rows + t * IN_ROW + wcol;
        dst[0] = x.x; dst[1] = x.y; dst[2] = x.z; dst[3] = x.w;
    }
    return tile_n;
}

// Per-thread summary (valid for any entry state) + inclusive block scan by composition.
// scan[] must hold 2*IN_THREADS entries.  Returns the thread's own summary; scan result of thread
// t is left in *incl_out; the buffer holding the results is returned through *res_buf.
__device__ __forceinline__ void summarise_and_scan(const uint32_t *rows, uint32_t tile_n, int tail_c, Summ *scan,
                                                   Summ **res_buf) {
    const int t = threadIdx.x;
    const int begin = t * IN_CHUNK;
    const int len = begin >= (int)tile_n ? 0 : min(IN_CHUNK, (int)tile_n - begin);
    const uint32_t *row = rows + t * IN_ROW;
    int next_c = -1;
    if (len > 0) {
        if (begin + len < (int)tile_n) next_c = (int)(rows[(t + 1) * IN_ROW] & 0xffu);
        else next_c = tail_c;
    }
    Summ s = summ_identity();
    if (len > 0) {
        const WalkOut h = walk_chunk<false>(row, len, next_c, HD, nullptr);
        const WalkOut q = walk_chunk<false>(row, len, next_c, SQ, nullptr);
        const int c0 = (int)(row[0] & 0xffu);
        const bool ls_is_header = c0 == '>';
        s.res[HD] = h.res; s.hdr[HD] = h.hdr; s.dirty[HD] = h.dirty;
        s.res[SQ] = q.res; s.hdr[SQ] = q.hdr; s.dirty[SQ] = q.dirty;
        const WalkOut &l = ls_is_header ? h : q;
        s.res[LS] = l.res; s.hdr[LS] = l.hdr + (ls_is_header ? 1u : 0u); s.dirty[LS] = l.dirty;
        s.out = (l.state) | (h.state << 2) | (q.state << 4);
    }
    // Hillis-Steele inclusive scan with the composition operator
    Summ *cur = scan, *nxt = scan + IN_THREADS;
    cur[t] = s;
    __syncthreads();
    for (int o = 1; o < IN_THREADS; o <<= 1) {
        Summ v = cur[t];
        if (t >= o) v = summ_compose(cur[t - o], v);
        nxt[t] = v;
        __syncthreads();
        Summ *tmp = cur; cur = nxt; nxt = tmp;
    }
    *res_buf = cur;
}

__device__ __forceinline__ int tile_tail_byte(const uint8_t *bytes, uint64_t n, uint64_t tile_base, uint32_t tile_n) {
    const uint64_t e = tile_base + tile_n;
    return e < n ? (int)bytes[e] : -1;
}

__global__ void __launch_bounds__(IN_THREADS) fasta_tile_summaries(const uint8_t *__restrict__ bytes, uint64_t n,
                                                                   Summ *__restrict__ tile_summ) {
    __shared__ uint32_t rows[IN_THREADS * IN_ROW + 4];
    __shared__ Summ scan[2 * IN_THREADS];
    const uint64_t tile_base = (uint64_t)blockIdx.x * IN_TILE;
    const uint32_t tile_n = load_tile(bytes, n, tile_base, rows);
    const int tail_c = tile_tail_byte(bytes, n, tile_base, tile_n);
    __syncthreads();
    Summ *res;
    summarise_and_scan(rows, tile_n, tail_c, scan, &res);
    if (threadIdx.x == IN_THREADS - 1) tile_summ[blockIdx.x] = res[IN_THREADS - 1];
}

struct TilePrefix { uint64_t res, hdr; uint32_t state, pad; };
struct IngestTotals { uint64_t res, hdr, dirty, state; };

// single CTA: exclusive scan of the tile summaries starting from state LS
__global__ void __launch_bounds__(IN_THREADS) fasta_tile_scan(const Summ *__restrict__ tile_summ, uint64_t n_tiles,
                                                              TilePrefix *__restrict__ prefix,
                                                              IngestTotals *__restrict__ totals) {
    __shared__ Summ scan[2 * IN_THREADS];
    const int t = threadIdx.x;
    uint32_t c_state = LS;
    uint64_t c_res = 0, c_hdr = 0, c_dirty = 0;
    for (uint64_t base = 0; base < n_tiles; base += IN_THREADS) {
        const uint64_t i = base + t;
        Summ *cur = scan, *nxt = scan + IN_THREADS;
        cur[t] = i < n_tiles ? tile_summ[i] : summ_identity();
        __syncthreads();
        for (int o = 1; o < IN_THREADS; o <<= 1) {
            Summ v = cur[t];
            if (t >= o) v = summ_compose(cur[t - o], v);
            nxt[t] = v;
            __syncthreads();
            Summ *tmp = cur; cur = nxt; nxt = tmp;
        }
        if (i < n_tiles) {
            TilePrefix p;
            if (t == 0) {
                p.res = c_res; p.hdr = c_hdr; p.state = c_state;
            } else {
                const Summ &e = cur[t - 1];
                p.res = c_res + e.res[c_state];
                p.hdr = c_hdr + e.hdr[c_state];
                p.state = (e.out >> (2 * c_state)) & 3u;
            }
            p.pad = 0;
            prefix[i] = p;
        }
        const Summ last = cur[IN_THREADS - 1];
        c_res += last.res[c_state];
        c_hdr += last.hdr[c_state];
        c_dirty += last.dirty[c_state];
        c_state = (last.out >> (2 * c_state)) & 3u;
        __syncthreads();
    }
    if (t == 0) {
        totals->res = c_res; totals->hdr = c_hdr; totals->dirty = c_dirty; totals->state = c_state;
    }
}

__global__ void __launch_bounds__(IN_THREADS) fasta_emit(const uint8_t *__restrict__ bytes, uint64_t n,
                                                         const TilePrefix *__restrict__ prefix,
                                                         unsigned long long *__restrict__ packed,
                                                         uint64_t res_store_base,   // residues already in the store
                                                         uint64_t *__restrict__ table, uint64_t table_cap) {
    __shared__ uint32_t rows[IN_THREADS * IN_ROW + 4];
    __shared__ Summ scan[2 * IN_THREADS];
    __shared__ unsigned long long words[IN_TILE / 32 + 2];
    const int t = threadIdx.x;
    const uint64_t tile_base = (uint64_t)blockIdx.x * IN_TILE;
    const uint32_t tile_n = load_tile(bytes, n, tile_base, rows);
    const int tail_c = tile_tail_byte(bytes, n, tile_base, tile_n);
    for (int i = t; i < IN_TILE / 32 + 2; i += IN_THREADS) words[i] = 0ULL;
    __syncthreads();
    Summ *res;
    summarise_and_scan(rows, tile_n, tail_c, scan, &res);
    const TilePrefix tp = prefix[blockIdx.x];
    uint32_t state = tp.state;
    uint32_t res_before = 0, hdr_before = 0;
    if (t > 0) {
        const Summ &e = res[t - 1];
        res_before = e.res[tp.state];
        hdr_before = e.hdr[tp.state];
        state = (e.out >> (2 * tp.state)) & 3u;
    }
    const uint32_t tile_res = res[IN_THREADS - 1].res[tp.state];
    const uint64_t gres0 = res_store_base + tp.res;     // global slot of the tile's first residue
    const uint32_t lead = (uint32_t)(gres0 & 31u);
    const int begin = t * IN_CHUNK;
    const int len = begin >= (int)tile_n ? 0 : min(IN_CHUNK, (int)tile_n - begin);
    if (len > 0) {
        int next_c = (begin + len < (int)tile_n) ? (int)(rows[(t + 1) * IN_ROW] & 0xffu) : tail_c;
        EmitCtx ctx;
        ctx.words = words;
        ctx.q = lead + res_before;
        ctx.res_global = tp.res + res_before;
        ctx.hdr_global = tp.hdr + hdr_before;
        ctx.byte_base = tile_base + begin;
        ctx.table = table;
        ctx.table_cap = table_cap;
        walk_chunk<true>(rows + t * IN_ROW, len, next_c, state, &ctx);
    }
    __syncthreads();
    if (tile_res == 0) return;
    const uint32_t nw = (lead + tile_res + 31) >> 5;
    const uint64_t gw0 = gres0 >> 5;
    for (uint32_t i = t; i < nw; i += IN_THREADS) {
        const unsigned long long v = words[i];
        if (i == 0 || i == nw - 1) {
            if (v) atomicOr(&packed[gw0 + i], v);   // word shared with the neighbouring tile
        } else {
            packed[gw0 + i] = v;
        }
    }
}

int ingest_fasta(pdg_graph *g, const uint8_t *d_bytes, uint64_t n_bytes, uint64_t *h_header_table,
                 uint64_t cap_headers, pdg_ingest_info *info) {
    cudaStream_t s = g->stream;
    pdg_ingest_info out{};
    out.n_bytes = n_bytes;
    out.residue_base = g->n_res;
    if (n_bytes == 0) {
        *info = out;
        out.appended = 1;
        *info = out;
        return PDG_OK;
    }
    const uint64_t n_tiles = (n_bytes + IN_TILE - 1) / IN_TILE;
    g->scratch0.reserve(n_tiles * sizeof(Summ), s);
    g->scratch1.reserve(n_tiles * sizeof(TilePrefix) + sizeof(IngestTotals), s);
    Summ *d_summ = g->scratch0.as<Summ>();
    TilePrefix *d_prefix = g->scratch1.as<TilePrefix>();
    IngestTotals *d_tot = reinterpret_cast<IngestTotals *>(g->scratch1.as<unsigned char>() + n_tiles * sizeof(TilePrefix));
    fasta_tile_summaries<<<(unsigned)n_tiles, IN_THREADS, 0, s>>>(d_bytes, n_bytes, d_summ);
    PDG_LAUNCHED(g);
    fasta_tile_scan<<<1, IN_THREADS, 0, s>>>(d_summ, n_tiles, d_prefix, d_tot);
    PDG_LAUNCHED(g);
    IngestTotals tot{};
    PDG_CUDA(cudaMemcpyAsync(&tot, d_tot, sizeof(tot), cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    out.n_headers = tot.hdr;
    out.n_residues = tot.res;
    out.n_dirty = tot.dirty;
    *info = out;
    if (tot.hdr > cap_headers) {
        set_error("pdg_ingest_fasta: %llu headers, table capacity %llu", (unsigned long long)tot.hdr,
                  (unsigned long long)cap_headers);
        return PDG_E_CAPACITY;
    }
    const bool append = tot.dirty == 0;
    // header table on the device (rows of 3), residues appended only for clean files
    g->small.reserve((tot.hdr * 3 + 3) * sizeof(uint64_t), s);
    uint64_t *d_table = g->small.as<uint64_t>();
    unsigned long long *d_packed = nullptr;
    if (append) {
        const uint64_t words = (g->n_res + tot.res + 31) / 32 + 2;
        g->packed.reserve(words * 8, s, /*keep=*/true, /*zero=*/true);
        d_packed = g->packed.as<unsigned long long>();
    } else {
        // still run pass C for the header table; residues go to a scratch word buffer
        const uint64_t words = (tot.res + 31) / 32 + 2;
        g->scratch2.reserve(words * 8, s);
        PDG_CUDA(cudaMemsetAsync(g->scratch2.p, 0, words * 8, s));
        d_packed = g->scratch2.as<unsigned long long>();
    }
    fasta_emit<<<(unsigned)n_tiles, IN_THREADS, 0, s>>>(d_bytes, n_bytes, d_prefix, d_packed,
                                                        append ? g->n_res : 0, d_table, tot.hdr);
    PDG_LAUNCHED(g);
    if (tot.hdr) {
        PDG_CUDA(cudaMemcpyAsync(h_header_table, d_table, tot.hdr * 3 * sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
    }
    PDG_CUDA(cudaStreamSynchronize(s));
    if (tot.hdr && tot.state == HD) h_header_table[3 * (tot.hdr - 1) + 1] = n_bytes;   // header ended by EOF
    if (append) {
        g->n_res += tot.res;
        out.appended = 1;
    }
    *info = out;
    return PDG_OK;
}

}  // namespace pdg
