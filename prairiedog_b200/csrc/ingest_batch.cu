// ingest_batch.cu -- K0/K1a, batched: many FASTA images -> 2-bit packed residues + header rows in
// three launches for the whole batch (pdg_ingest_fasta_batch), word-parallel instead of byte-serial.
//
// Restates Kmers._load (prairiedog/kmers.py:51-77) for every file of the batch.  A thread owns 64
// bytes and turns them into 64-bit masks with SIMD-in-register byte tests (N = '\n', G = '>',
// V = one of ACGT) plus the 2-bit codes of all 64 bytes.  Line structure follows from the masks:
//   L = (N << 1) | (byte 0 starts a line)        line starts
//   H = L & G                                    starts of '>' lines          (kmers.py:64-71)
//   inH = (N - H) & ~N                           bytes inside '>' lines: the borrow of the
//                                                subtraction runs from each H bit to its newline
// The only thing a chunk needs from its left is one bit (does it begin inside a '>' line), and a
// chunk either fixes that bit for its right neighbour (it contains a line start / newline) or
// passes it through, so the scan over chunks, tiles and files is a "last writer wins" propagation
// done with warp ballots.
//   pass A  fb_summ_kernel   1 B/byte read                 -> one summary per 16 KiB tile
//   pass B  fb_scan_kernel   one CTA per file              -> tile prefixes + file totals
//   pass C  fb_emit_kernel   1 B/byte read, 0.25 B/residue -> packed store + header rows
//           fb_hash_kernel   a 64-bit hash per header line (duplicate-label detection on the host)
// This path handles PURE files: every byte is inside a '>' line, is one of ACGT or is '\n', and
// the file holds no '\r'.  If any file of the batch is not pure (CRLF line ends, trailing blanks,
// N / IUPAC / lower case) the whole batch goes through the exact byte automaton of ingest.cu, file
// by file, so results never depend on which path ran (tests/test_gpu_ingest_batch.py).
// With PDG_IN_HOST the images are host buffers: they are copied to a device staging area on a
// second stream in groups, and pass A of a group overlaps the copy of the next one.
#include "gindex.cuh"
#include "msd3.cuh"
#include <algorithm>
#include <stdlib.h>

namespace pdg {

static constexpr int FB_THREADS = 256;
static constexpr int FB_WARPS = FB_THREADS / 32;
static constexpr int FB_CHUNK = 64;                        // bytes per thread
static constexpr int FB_TILE = FB_THREADS * FB_CHUNK;      // 16 KiB per CTA
static constexpr int FB_WORDS = FB_TILE / 32 + 2;          // packed words a tile can touch
static constexpr uint64_t FB_GROUP_BYTES = 32ull << 20;    // host mode: copy / pass A granularity

struct FbFile {
    const uint8_t *bytes;    // device image
    uint64_t n_bytes;
    uint64_t tile0;          // first tile of the file inside the batch
    uint64_t res_base;       // store slot of the file's first residue (pass C)
    uint64_t hdr_base;       // first header row of the file in the batch table (pass C)
};

struct FbTileSumm {
    uint32_t res[2];         // residues of the tile when it begins outside / inside a '>' line
    uint32_t hdr;            // '>' lines started in the tile
    uint32_t flags;          // bit0 exit state is fixed, bit1 its value, bit2/3 impure for entry 0/1
};
struct FbTilePrefix { uint64_t res; uint32_t hdr; uint32_t state; };   // file relative, before the tile
struct FbFileTotals { uint64_t res, hdr; uint32_t impure, state; };

struct FbChunk {
    uint64_t N, G, V;        // bit i = byte i of the chunk
    uint64_t c0, c1;         // 2-bit codes of bytes 0..31 / 32..63, first byte most significant
    uint32_t len;
    bool cr, ls0;
};

__device__ __forceinline__ uint32_t fb_zero_bytes(uint32_t x) {   // 0x80 in every byte of x that is 0
    return ~(((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) & 0x80808080u;
}
__device__ __forceinline__ uint32_t fb_gather(uint32_t z) {       // bit i = flag of byte i
    return ((z >> 7) * 0x01020408u) >> 24;
}

// last file whose tile0 <= tile (files[n_files] is a sentinel)
__device__ __forceinline__ uint32_t fb_file_of(const FbFile *__restrict__ files, uint32_t n_files, uint64_t tile) {
    uint32_t lo = 0, hi = n_files;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (files[mid].tile0 <= tile) lo = mid + 1; else hi = mid;
    }
    return lo - 1;
}

template <bool CODES>
__device__ __forceinline__ void fb_load_chunk(const uint8_t *__restrict__ bytes, uint64_t n, uint64_t off, FbChunk &c) {
    c.len = off >= n ? 0u : (uint32_t)min((uint64_t)FB_CHUNK, n - off);
    uint32_t w[16];
    if (c.len == FB_CHUNK && ((reinterpret_cast<uintptr_t>(bytes) + off) & 15) == 0) {
        const uint4 *p = reinterpret_cast<const uint4 *>(bytes + off);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint4 x = __ldg(p + i);
            w[4 * i] = x.x; w[4 * i + 1] = x.y; w[4 * i + 2] = x.z; w[4 * i + 3] = x.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) {   // unaligned image or the tail of a file (static indices: w stays in registers)
            uint32_t x = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if ((uint32_t)(4 * i + j) < c.len) x |= (uint32_t)bytes[off + 4 * i + j] << (8 * j);
            w[i] = x;
        }
    }
    c.ls0 = c.len > 0 && (off == 0 || bytes[off - 1] == '\n');
    uint32_t n_lo = 0, n_hi = 0, g_lo = 0, g_hi = 0, v_lo = 0, v_hi = 0, crz = 0;
    uint32_t cw[4] = {0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const uint32_t x = w[i];
        const uint32_t nl = fb_gather(fb_zero_bytes(x ^ 0x0a0a0a0au));
        const uint32_t gt = fb_gather(fb_zero_bytes(x ^ 0x3e3e3e3eu));
        const uint32_t t = (x >> 1) & 0x03030303u;                     // A0 C1 T2 G3 for the four letters
        const uint32_t is_t = (t >> 1) & ~t & 0x01010101u;
        const uint32_t expect = (0x41414141u + (x & 0x06060606u)) ^ (is_t * 0x11u);   // 41 43 54 47
        const uint32_t va = fb_gather(fb_zero_bytes(x ^ expect));
        crz |= fb_zero_bytes(x ^ 0x0d0d0d0du);
        const int sh = 4 * (i & 7);
        if (i < 8) { n_lo |= nl << sh; g_lo |= gt << sh; v_lo |= va << sh; }
        else { n_hi |= nl << sh; g_hi |= gt << sh; v_hi |= va << sh; }
        if (CODES) {
            // A=0 T=1 C=2 G=3 (kmers.py:23 product order): swap the two bits of t, pack 4 bytes -> 8 bits
            const uint32_t y = ((t & 0x01010101u) << 1) | ((t >> 1) & 0x01010101u);
            cw[i >> 2] = (cw[i >> 2] << 8) | ((y * 0x40100401u) >> 24);
        }
    }
    c.N = ((uint64_t)n_hi << 32) | n_lo;
    c.G = ((uint64_t)g_hi << 32) | g_lo;
    c.V = ((uint64_t)v_hi << 32) | v_lo;
    c.cr = crz != 0;
    c.c0 = ((uint64_t)cw[0] << 32) | cw[1];
    c.c1 = ((uint64_t)cw[2] << 32) | cw[3];
}

// "Last writer wins" over the threads of a CTA.  set/val: this thread fixes the state seen by its
// right neighbour.  Out: fixed/ent = a thread to the left fixed this thread's entry state;
// returns the CTA summary (bit0 some thread fixes the exit state, bit1 its value).
__device__ __forceinline__ uint32_t fb_propagate(bool set, bool val, uint32_t *s_warp, bool &fixed, uint32_t &ent) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t sm = __ballot_sync(0xffffffffu, set), vm = __ballot_sync(0xffffffffu, val);
    const uint32_t below = sm & ((1u << lane) - 1u);
    fixed = below != 0;
    ent = fixed ? (vm >> (31 - __clz(below))) & 1u : 0u;
    __syncthreads();   // s_warp may still be read by a previous call
    if (lane == 0) s_warp[w] = sm ? (1u | (((vm >> (31 - __clz(sm))) & 1u) << 1)) : 0u;
    __syncthreads();
    if (!fixed)
        for (int ww = w - 1; ww >= 0; --ww)
            if (s_warp[ww] & 1u) { fixed = true; ent = s_warp[ww] >> 1; break; }
    uint32_t kind = 0;
    for (int ww = FB_WARPS - 1; ww >= 0; --ww)
        if (s_warp[ww] & 1u) { kind = s_warp[ww]; break; }
    return kind;
}

__device__ __forceinline__ uint64_t fb_len_mask(uint32_t len) { return len >= 64 ? ~0ULL : ((1ULL << len) - 1ULL); }

// ---- pass A ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(FB_THREADS) fb_summ_kernel(const FbFile *__restrict__ files, uint32_t n_files,
                                                             uint64_t tile_begin, FbTileSumm *__restrict__ out) {
    __shared__ uint32_t s_warp[FB_WARPS];
    __shared__ uint32_t s_red[FB_WARPS][4];
    const uint64_t tile = tile_begin + blockIdx.x;
    const uint32_t f = fb_file_of(files, n_files, tile);
    const FbFile fi = files[f];
    const uint64_t off = (tile - fi.tile0) * FB_TILE + (uint64_t)threadIdx.x * FB_CHUNK;
    FbChunk c;
    fb_load_chunk<false>(fi.bytes, fi.n_bytes, off, c);
    const uint64_t L = (c.N << 1) | (c.ls0 ? 1ULL : 0ULL);
    const uint64_t H = L & c.G, H1 = H | (c.ls0 ? 0ULL : 1ULL);
    const bool exit0 = c.N < H, exit1 = c.N < H1;
    const uint64_t in0 = (c.N - H) & ~c.N, in1 = (c.N - H1) & ~c.N;
    const uint64_t body = fb_len_mask(c.len) & ~c.N;
    const uint32_t res0 = __popcll(c.V & ~in0), res1 = __popcll(c.V & ~in1);
    const bool imp0 = c.cr || (body & ~c.V & ~in0) != 0, imp1 = c.cr || (body & ~c.V & ~in1) != 0;
    bool fixed;
    uint32_t ent;
    const uint32_t kind = fb_propagate(exit0 == exit1, exit0, s_warp, fixed, ent);
    uint32_t a0 = fixed ? (ent ? res1 : res0) : res0, a1 = fixed ? (ent ? res1 : res0) : res1;
    uint32_t hd = __popcll(H);
    uint32_t im = (fixed ? (ent ? imp1 : imp0) : imp0) ? 1u : 0u;
    im |= ((fixed ? (ent ? imp1 : imp0) : imp1) ? 2u : 0u);
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        a0 += __shfl_xor_sync(0xffffffffu, a0, o);
        a1 += __shfl_xor_sync(0xffffffffu, a1, o);
        hd += __shfl_xor_sync(0xffffffffu, hd, o);
        im |= __shfl_xor_sync(0xffffffffu, im, o);
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { s_red[w][0] = a0; s_red[w][1] = a1; s_red[w][2] = hd; s_red[w][3] = im; }
    __syncthreads();
    if (threadIdx.x == 0) {
        FbTileSumm s{{0u, 0u}, 0u, 0u};
        uint32_t imp = 0;
        for (int ww = 0; ww < FB_WARPS; ++ww) {
            s.res[0] += s_red[ww][0]; s.res[1] += s_red[ww][1]; s.hdr += s_red[ww][2]; imp |= s_red[ww][3];
        }
        s.flags = kind | (imp << 2);
        out[tile] = s;
    }
}

// ---- pass B: one CTA per file -------------------------------------------------------------------
__global__ void __launch_bounds__(FB_THREADS) fb_scan_kernel(const FbFile *__restrict__ files, uint32_t f_begin,
                                                             const FbTileSumm *__restrict__ summ,
                                                             FbTilePrefix *__restrict__ prefix,
                                                             FbFileTotals *__restrict__ totals) {
    __shared__ uint32_t s_warp[FB_WARPS];
    __shared__ uint32_t s_imp;
    const uint32_t f = f_begin + blockIdx.x;
    const uint64_t t0 = files[f].tile0, nt = files[f + 1].tile0 - t0;
    uint64_t c_res = 0, c_hdr = 0;
    uint32_t c_state = 0;
    if (threadIdx.x == 0) s_imp = 0;
    __syncthreads();
    for (uint64_t base = 0; base < nt; base += FB_THREADS) {
        const uint64_t i = base + threadIdx.x;
        FbTileSumm s{{0u, 0u}, 0u, 0u};
        if (i < nt) s = summ[t0 + i];
        bool fixed;
        uint32_t ent;
        const uint32_t kind = fb_propagate((s.flags & 1u) != 0, (s.flags & 2u) != 0, s_warp, fixed, ent);
        const uint32_t e = fixed ? ent : c_state;
        uint64_t tot_r, tot_h;
        const uint64_t ex_r = block_excl_scan((uint64_t)s.res[e], &tot_r);
        const uint64_t ex_h = block_excl_scan((uint64_t)s.hdr, &tot_h);
        if (i < nt) {
            prefix[t0 + i] = FbTilePrefix{c_res + ex_r, (uint32_t)(c_hdr + ex_h), e};
            if ((s.flags >> (2 + e)) & 1u) s_imp = 1u;
        }
        c_res += tot_r;
        c_hdr += tot_h;
        if (kind & 1u) c_state = kind >> 1;
    }
    __syncthreads();
    if (threadIdx.x == 0) totals[f] = FbFileTotals{c_res, c_hdr, s_imp, c_state};
}

// ---- pass C ------------------------------------------------------------------------------------
// ORs `cnt` residues (2 bits each, right aligned in `bits`) into the tile's word buffer at slot q
__device__ __forceinline__ void fb_put(unsigned long long *words, uint32_t q, uint64_t bits, uint32_t cnt) {
    const uint32_t wd = q >> 5, o = q & 31u, space = 32u - o;
    if (cnt <= space) {
        atomicOr(&words[wd], (unsigned long long)(bits << (2u * (space - cnt))));
    } else {
        const uint32_t rest = cnt - space;
        atomicOr(&words[wd], (unsigned long long)(bits >> (2u * rest)));
        atomicOr(&words[wd + 1], (unsigned long long)(bits << (64u - 2u * rest)));
    }
}
// codes of bytes [s, s+cnt) of one 32-byte half (cnt >= 1, s + cnt <= 32), right aligned
__device__ __forceinline__ uint64_t fb_slice(uint64_t half, uint32_t s, uint32_t cnt) {
    return (half << (2u * s)) >> (64u - 2u * cnt);
}

__global__ void __launch_bounds__(FB_THREADS) fb_emit_kernel(const FbFile *__restrict__ files, uint32_t n_files,
                                                             uint64_t tile_begin,
                                                             const FbTilePrefix *__restrict__ prefix,
                                                             unsigned long long *__restrict__ packed,
                                                             uint64_t *__restrict__ table) {
    __shared__ uint32_t s_warp[FB_WARPS];
    __shared__ unsigned long long words[FB_WORDS];
    const uint64_t tile = tile_begin + blockIdx.x;
    const uint32_t f = fb_file_of(files, n_files, tile);
    const FbFile fi = files[f];
    const FbTilePrefix tp = prefix[tile];
    const uint64_t off = (tile - fi.tile0) * FB_TILE + (uint64_t)threadIdx.x * FB_CHUNK;
    for (int i = threadIdx.x; i < FB_WORDS; i += FB_THREADS) words[i] = 0ULL;
    FbChunk c;
    fb_load_chunk<true>(fi.bytes, fi.n_bytes, off, c);
    const uint64_t L = (c.N << 1) | (c.ls0 ? 1ULL : 0ULL);
    const uint64_t H = L & c.G, H1 = H | (c.ls0 ? 0ULL : 1ULL);
    const bool exit0 = c.N < H, exit1 = c.N < H1;
    bool fixed;
    uint32_t ent;
    fb_propagate(exit0 == exit1, exit0, s_warp, fixed, ent);
    const bool cont = (fixed ? ent : tp.state) != 0 && !c.ls0 && c.len > 0;   // chunk begins inside a '>' line
    const uint64_t inH = (c.N - (cont ? H1 : H)) & ~c.N;
    uint64_t S = c.V & ~inH;
    uint64_t tile_res, tile_hdr;
    const uint32_t res_before = (uint32_t)block_excl_scan((uint64_t)__popcll(S), &tile_res);
    const uint32_t hdr_before = (uint32_t)block_excl_scan((uint64_t)__popcll(H), &tile_hdr);
    const uint64_t gres0 = fi.res_base + tp.res;       // store slot of the tile's first residue
    const uint32_t lead = (uint32_t)(gres0 & 31u);
    // header rows {byte_begin, byte_end, residue_index, hash}
    {
        const uint64_t row0 = fi.hdr_base + tp.hdr + hdr_before;
        if (cont && c.N) table[4 * (row0 - 1) + 1] = off + (uint32_t)(__ffsll((long long)c.N) - 1);
        uint64_t hr = H;
        uint32_t i = 0;
        while (hr) {
            const uint32_t a = (uint32_t)(__ffsll((long long)hr) - 1);
            hr &= hr - 1;
            uint64_t *row = table + 4 * (row0 + i);
            row[0] = off + a;
            row[2] = tp.res + res_before + __popcll(S & ((1ULL << a) - 1ULL));
            const uint64_t above = c.N & ~((2ULL << a) - 1ULL);
            if (above) row[1] = off + (uint32_t)(__ffsll((long long)above) - 1);
            ++i;
        }
    }
    // residues: runs of selected bytes -> 2-bit codes at slot lead + res_before
    uint32_t q = lead + res_before;
    while (S) {
        const uint32_t a = (uint32_t)(__ffsll((long long)S) - 1);
        const uint64_t t = ~(S >> a);
        const uint32_t run = t ? (uint32_t)(__ffsll((long long)t) - 1) : 64u - a;
        S = (a + run >= 64u) ? 0ULL : (S >> (a + run)) << (a + run);
        const uint32_t e = a + run;
        if (a < 32u) {
            const uint32_t cnt = min(e, 32u) - a;
            fb_put(words, q, fb_slice(c.c0, a, cnt), cnt);
            q += cnt;
        }
        if (e > 32u) {
            const uint32_t s = max(a, 32u) - 32u, cnt = e - 32u - s;
            fb_put(words, q, fb_slice(c.c1, s, cnt), cnt);
            q += cnt;
        }
    }
    __syncthreads();
    if (tile_res == 0) return;
    const uint32_t nw = (lead + (uint32_t)tile_res + 31u) >> 5;
    const uint64_t gw0 = gres0 >> 5;
    for (uint32_t i = threadIdx.x; i < nw; i += FB_THREADS) {
        const unsigned long long v = words[i];
        if (i == 0 || i == nw - 1) {
            if (v) atomicOr(&packed[gw0 + i], v);   // word shared with the neighbouring tile / file
        } else {
            packed[gw0 + i] = v;
        }
    }
}

// One thread per header row: closes a header ended by end of file and hashes the right-stripped
// line (FNV-1a).  Hash 0 = "compare the bytes on the host" (line ends in a non-ASCII byte).
// last file of [f0, f1) whose hdr_base <= r (files without rows share the base of the next one)
__device__ __forceinline__ uint32_t fb_file_of_row(const FbFile *__restrict__ files, uint32_t f0, uint32_t f1, uint64_t r) {
    uint32_t lo = f0, hi = f1;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (files[mid].hdr_base <= r) lo = mid + 1; else hi = mid;
    }
    return lo - 1;
}

__global__ void fb_hash_kernel(const FbFile *__restrict__ files, uint32_t f0, uint32_t f1,
                               const FbFileTotals *__restrict__ totals, uint64_t row_begin, uint64_t row_end,
                               uint64_t *__restrict__ table) {
    const uint64_t r = row_begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= row_end) return;
    const uint32_t f = fb_file_of_row(files, f0, f1, r);
    const FbFile fi = files[f];
    uint64_t *row = table + 4 * r;
    if (r + 1 == fi.hdr_base + totals[f].hdr && totals[f].state == 1u) row[1] = fi.n_bytes;
    const uint64_t b = row[0];
    uint64_t e = row[1];
    while (e > b) {
        const uint8_t ch = fi.bytes[e - 1];
        if ((ch >= 9 && ch <= 13) || (ch >= 28 && ch <= 32)) --e; else break;
    }
    uint64_t h = 0xcbf29ce484222325ULL;
    for (uint64_t i = b; i < e; ++i) h = (h ^ fi.bytes[i]) * 0x100000001b3ULL;
    if (h == 0) h = 1;
    if (e > b && fi.bytes[e - 1] >= 0x80) h = 0;
    row[3] = h;
}

// Header rows -> the contig table the reference's loader would build from a well-formed file
// (kmers.py:61-75): row r covers residues [res_base + residue_index(r), start of the next row or
// the end of the file).  Used by the pipelined build; the host checks later that the caller
// declared exactly these contigs.
__global__ void fb_contigs_kernel(const FbFile *__restrict__ files, uint32_t f0, uint32_t f1,
                                  const FbFileTotals *__restrict__ totals, uint64_t row_begin, uint64_t row_end,
                                  const uint64_t *__restrict__ table, int k, uint64_t *__restrict__ ct_begin,
                                  uint64_t *__restrict__ ct_end) {
    const uint64_t r = row_begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= row_end) return;
    const uint32_t f = fb_file_of_row(files, f0, f1, r);
    const FbFile fi = files[f];
    const bool last = r + 1 == fi.hdr_base + totals[f].hdr;
    const uint64_t b = fi.res_base + table[4 * r + 2];
    const uint64_t e = fi.res_base + (last ? totals[f].res : table[4 * (r + 1) + 2]);
    ct_begin[r] = b;
    // a contig with len <= k yields nothing (kmers.py:110-117) -- not even the one window of len == k,
    // and pdg_add_contigs drops it: keep it in the table (rows stay aligned) but empty
    ct_end[r] = e - b > (uint64_t)k ? e : b;
}

// ---- host -----------------------------------------------------------------------------------------
static int batch_fallback(pdg_graph *g, const std::vector<const uint8_t *> &dev, const uint64_t *n_bytes,
                          uint32_t n_files, uint64_t *h_table, pdg_ingest_info *infos) {
    std::vector<uint64_t> rows3;
    uint64_t row = 0;
    for (uint32_t f = 0; f < n_files; ++f) {
        const uint64_t cap = infos[f].n_headers;
        rows3.assign(3 * cap + 3, 0);
        const int rc = ingest_fasta(g, dev[f], n_bytes[f], rows3.data(), cap, &infos[f]);
        if (rc != PDG_OK) return rc;
        for (uint64_t i = 0; i < infos[f].n_headers; ++i) {
            h_table[4 * (row + i) + 0] = rows3[3 * i + 0];
            h_table[4 * (row + i) + 1] = rows3[3 * i + 1];
            h_table[4 * (row + i) + 2] = rows3[3 * i + 2];
            h_table[4 * (row + i) + 3] = 0;   // hash unknown: the host compares bytes
        }
        row += infos[f].n_headers;
    }
    return PDG_OK;
}

// ---- pipelined build (host images, k <= 11, empty graph) -------------------------------------------
// The copy of a 500 MB batch takes ~10 ms over PCIe 5 while parsing it takes < 1 ms, and the index
// sort cannot finish before the last byte arrived.  So the batch is cut into copy groups and every
// group is taken as far as it can go while the next ones are still in flight:
//   copy stream : H2D group 0 | H2D group 1 | H2D group 2 | ...
//   aux stream  :      pass A+B g0 | pass A+B g1 | ...                (needs only the group's bytes)
//   main stream :        emit g0, contig table g0, hist + pass 1 + pass 2 g0 | emit g1, ...   (msd3_stage_group)
// After the last group only pass 3 and the row table are left for pdg_finalize (msd3_finish).  The
// library has to assume which contigs the caller will declare (pdg_add_contigs comes later): the ones
// a well-formed file yields, one per header row.  pdg_finalize checks the declaration against that
// assumption and falls back to the one-shot build when it differs.
struct FbGroup { uint32_t f0, f1; uint64_t tb, te; };

static int ingest_pipelined(pdg_graph *g, const uint8_t *const *ptrs, const uint64_t *n_bytes, uint32_t n_files,
                            uint64_t *h_table, uint64_t cap_headers, pdg_ingest_info *infos, std::vector<FbFile> &hf,
                            const std::vector<const uint8_t *> &dev, uint64_t tiles, uint64_t group_bytes, bool host_src) {
    cudaStream_t s = g->stream;
    const int k = g->k;
    // ---- groups ----
    std::vector<FbGroup> groups;
    uint64_t total_bytes = 0;
    // (groups that taper off towards the end were tried: the fixed cost of a group -- a host round trip
    // for the totals and ~12 launches -- outweighs the shorter tail; 16.1 vs 15.1 ms per config-3 step)
    for (uint32_t f0 = 0; f0 < n_files;) {
        uint32_t f1 = f0;
        uint64_t bytes = 0;
        while (f1 < n_files && (f1 == f0 || bytes + n_bytes[f1] <= group_bytes)) bytes += n_bytes[f1++];
        groups.push_back(FbGroup{f0, f1, hf[f0].tile0, hf[f1].tile0});
        total_bytes += bytes;
        f0 = f1;
    }
    // (also tried: a last group of a single file -- 14.26 vs 13.94 ms; a group's fixed cost again)
    const uint32_t G = (uint32_t)groups.size();
    // ---- every buffer up front (a reallocation in the middle would stall the pipeline) ----
    const uint64_t words = (total_bytes + 31) / 32 + 2;          // residues <= bytes
    g->packed.reserve(words * 8, s, /*keep=*/false, /*zero=*/true);
    g->small.reserve((cap_headers * 4 + 8) * sizeof(uint64_t), s);
    g->pre_ct.reserve((2 * cap_headers + 4) * sizeof(uint64_t), s);
    msd3_reserve(g, 0, std::max<uint64_t>(total_bytes, 32), G);
    if (!g->aux_stream) PDG_CUDA(cudaStreamCreateWithFlags(&g->aux_stream, cudaStreamNonBlocking));
    while (g->fb_events.size() < 2 * (size_t)G + 1) {
        cudaEvent_t ev;
        PDG_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        g->fb_events.push_back(ev);
    }
    FbFile *d_files = g->fb_files.as<FbFile>();
    FbTileSumm *d_summ = g->fb_summ.as<FbTileSumm>();
    FbTilePrefix *d_prefix = g->fb_prefix.as<FbTilePrefix>();
    FbFileTotals *d_tot = g->fb_totals.as<FbFileTotals>();
    uint64_t *d_table = g->small.as<uint64_t>();
    uint64_t *ct_begin = g->pre_ct.as<uint64_t>(), *ct_end = ct_begin + cap_headers + 2;
    // everything queued on s so far (file table upload, reserves, staging area readers) precedes the copies
    cudaEvent_t ev_ready = g->fb_events[2 * G];
    PDG_CUDA(cudaEventRecord(ev_ready, s));
    PDG_CUDA(cudaStreamWaitEvent(g->aux_stream, ev_ready, 0));
    if (host_src) {
        PDG_CUDA(cudaStreamWaitEvent(g->copy_stream, ev_ready, 0));
        for (uint32_t gi = 0; gi < G; ++gi) {
            for (uint32_t f = groups[gi].f0; f < groups[gi].f1; ++f)
                if (n_bytes[f])
                    PDG_CUDA(cudaMemcpyAsync(const_cast<uint8_t *>(dev[f]), ptrs[f], n_bytes[f], cudaMemcpyHostToDevice,
                                             g->copy_stream));
            PDG_CUDA(cudaEventRecord(g->fb_events[2 * gi], g->copy_stream));
        }
        PDG_CUDA(cudaEventRecord(g->i_events[2], g->copy_stream));   // last byte on the device
    }
    const bool time_kernels = !host_src && G == 1;   // device images: one group, its launches run alone
    std::vector<FbFileTotals> tot(n_files);
    uint64_t n_rows = 0, n_res = 0;
    bool pure = true, fits = true;
    for (uint32_t gi = 0; gi < G; ++gi) {
        const FbGroup &gr = groups[gi];
        const uint32_t nf = gr.f1 - gr.f0;
        // pass A + B of the group as soon as its bytes are on the device
        if (host_src) PDG_CUDA(cudaStreamWaitEvent(g->aux_stream, g->fb_events[2 * gi], 0));
        if (time_kernels) PDG_CUDA(cudaEventRecord(g->i_events[6], g->aux_stream));
        if (gr.te > gr.tb) {
            fb_summ_kernel<<<(unsigned)(gr.te - gr.tb), FB_THREADS, 0, g->aux_stream>>>(d_files, n_files, gr.tb, d_summ);
            PDG_LAUNCHED(g);
        }
        if (time_kernels) PDG_CUDA(cudaEventRecord(g->i_events[7], g->aux_stream));
        fb_scan_kernel<<<nf, FB_THREADS, 0, g->aux_stream>>>(d_files, gr.f0, d_summ, d_prefix, d_tot);
        PDG_LAUNCHED(g);
        PDG_CUDA(cudaMemcpyAsync(tot.data() + gr.f0, d_tot + gr.f0, nf * sizeof(FbFileTotals), cudaMemcpyDeviceToHost,
                                 g->aux_stream));
        PDG_CUDA(cudaEventRecord(g->fb_events[2 * gi + 1], g->aux_stream));
        PDG_CUDA(cudaEventSynchronize(g->fb_events[2 * gi + 1]));
        const uint64_t row_begin = n_rows, res_begin = n_res;
        for (uint32_t f = gr.f0; f < gr.f1; ++f) {
            pdg_ingest_info o{};
            o.n_bytes = n_bytes[f];
            o.n_headers = tot[f].hdr;
            o.n_residues = tot[f].res;
            o.residue_base = g->n_res + n_res;
            o.appended = 1;
            infos[f] = o;
            hf[f].res_base = g->n_res + n_res;
            hf[f].hdr_base = n_rows;
            n_rows += tot[f].hdr;
            n_res += tot[f].res;
            pure = pure && tot[f].impure == 0;
        }
        fits = fits && n_rows <= cap_headers;
        if (!pure || !fits) continue;   // keep collecting the totals of every file, stage nothing more
        PDG_CUDA(cudaMemcpyAsync(d_files + gr.f0, hf.data() + gr.f0, nf * sizeof(FbFile), cudaMemcpyHostToDevice, s));
        if (time_kernels) PDG_CUDA(cudaEventRecord(g->i_events[3], s));
        if (gr.te > gr.tb) {
            fb_emit_kernel<<<(unsigned)(gr.te - gr.tb), FB_THREADS, 0, s>>>(d_files, n_files, gr.tb, d_prefix,
                                                                           g->packed.as<unsigned long long>(), d_table);
            PDG_LAUNCHED(g);
        }
        if (time_kernels) PDG_CUDA(cudaEventRecord(g->i_events[4], s));
        const uint64_t rows = n_rows - row_begin;
        if (rows) {
            const unsigned grid = (unsigned)((rows + 127) / 128);
            fb_hash_kernel<<<grid, 128, 0, s>>>(d_files, gr.f0, gr.f1, d_tot, row_begin, n_rows, d_table);
            PDG_LAUNCHED(g);
            fb_contigs_kernel<<<grid, 128, 0, s>>>(d_files, gr.f0, gr.f1, d_tot, row_begin, n_rows, d_table, k, ct_begin, ct_end);
            PDG_LAUNCHED(g);
        }
        if (n_res > res_begin && rows) {
            ContigTable ct{};
            ct.begin = ct_begin + row_begin;
            ct.end = ct_end + row_begin;
            ct.kbase = nullptr;
            ct.type = nullptr;
            ct.n = rows;
            msd3_stage_group(g, ct, g->n_res + res_begin, g->n_res + n_res);
        }
    }
    hf[n_files].hdr_base = n_rows;
    if (!fits || !pure) {
        // nothing may stay behind: the store beyond n_res has to be zero for the next append
        PDG_CUDA(cudaStreamSynchronize(g->aux_stream));
        PDG_CUDA(cudaMemsetAsync(g->packed.p, 0, g->packed.cap, s));
        g->m3_groups.clear();
        if (!fits) {
            set_error("pdg_ingest_fasta_batch: %llu headers, table capacity %llu", (unsigned long long)n_rows,
                      (unsigned long long)cap_headers);
            PDG_CUDA(cudaStreamSynchronize(s));
            return PDG_E_CAPACITY;
        }
        return batch_fallback(g, dev, n_bytes, n_files, h_table, infos);
    }
    // the header table goes back on the aux stream so that the main stream can carry on with pass 3
    // (speculatively: pdg_finalize redoes the build if the caller declares other contigs) while the
    // caller does its bookkeeping
    PDG_CUDA(cudaEventRecord(g->i_events[5], s));                 // end of the ingest proper
    PDG_CUDA(cudaStreamWaitEvent(g->aux_stream, g->i_events[5], 0));
    if (n_rows)
        PDG_CUDA(cudaMemcpyAsync(h_table, d_table, n_rows * 4 * sizeof(uint64_t), cudaMemcpyDeviceToHost, g->aux_stream));
    if (!g->m3_groups.empty()) msd3_finish_launch(g, /*rows=*/2 * k <= kMsd3MaxBits);
    PDG_CUDA(cudaStreamSynchronize(g->aux_stream));
    // the contigs this build assumed (host copy, checked by pdg_finalize)
    g->pre_begin.clear();
    g->pre_len.clear();
    for (uint32_t f = 0; f < n_files; ++f)
        for (uint64_t r = hf[f].hdr_base; r < hf[f].hdr_base + tot[f].hdr; ++r) {
            const uint64_t b = hf[f].res_base + h_table[4 * r + 2];
            const uint64_t e = hf[f].res_base + (r + 1 == hf[f].hdr_base + tot[f].hdr ? tot[f].res : h_table[4 * (r + 1) + 2]);
            if (e - b > (uint64_t)k) {
                g->pre_begin.push_back(b);
                g->pre_len.push_back(e - b);
            }
        }
    g->pre_valid = !g->m3_groups.empty();
    g->n_res += n_res;
    return PDG_OK;
}

int ingest_fasta_batch(pdg_graph *g, const uint8_t *const *ptrs, const uint64_t *n_bytes, uint32_t n_files,
                       int flags, uint64_t *h_table, uint64_t cap_headers, pdg_ingest_info *infos) {
    cudaStream_t s = g->stream;
    if (n_files == 0) return PDG_OK;
    const bool host_src = (flags & PDG_IN_HOST) != 0;
    // ---- file table ----
    std::vector<FbFile> hf(n_files + 1);
    std::vector<const uint8_t *> dev(n_files);
    uint64_t tiles = 0, stage_bytes = 0;
    std::vector<uint64_t> stage_off(n_files);
    for (uint32_t f = 0; f < n_files; ++f) {
        if (n_bytes[f] && !ptrs[f]) {
            set_error("pdg_ingest_fasta_batch: file %u has a null image", f);
            return PDG_E_INVALID;
        }
        stage_off[f] = stage_bytes;
        stage_bytes += (n_bytes[f] + 255) & ~255ULL;
        hf[f].n_bytes = n_bytes[f];
        hf[f].tile0 = tiles;
        hf[f].res_base = hf[f].hdr_base = 0;
        tiles += (n_bytes[f] + FB_TILE - 1) / FB_TILE;
    }
    hf[n_files] = FbFile{nullptr, 0, tiles, 0, 0};
    if (tiles >= (1ULL << 31)) {
        set_error("pdg_ingest_fasta_batch: batch of %llu tiles is too large", (unsigned long long)tiles);
        return PDG_E_INVALID;
    }
    if (host_src) {
        g->stage.reserve(std::max<uint64_t>(stage_bytes, 256), s);
        if (!g->copy_stream) PDG_CUDA(cudaStreamCreateWithFlags(&g->copy_stream, cudaStreamNonBlocking));
    }
    for (uint32_t f = 0; f < n_files; ++f)
        dev[f] = host_src ? g->stage.as<uint8_t>() + stage_off[f] : ptrs[f];
    for (uint32_t f = 0; f < n_files; ++f) hf[f].bytes = dev[f];
    g->fb_files.reserve((n_files + 1) * sizeof(FbFile), s);
    g->fb_summ.reserve(std::max<uint64_t>(tiles, 1) * sizeof(FbTileSumm), s);
    g->fb_prefix.reserve(std::max<uint64_t>(tiles, 1) * sizeof(FbTilePrefix), s);
    g->fb_totals.reserve(n_files * sizeof(FbFileTotals), s);
    FbFile *d_files = g->fb_files.as<FbFile>();
    FbTileSumm *d_summ = g->fb_summ.as<FbTileSumm>();
    FbTilePrefix *d_prefix = g->fb_prefix.as<FbTilePrefix>();
    FbFileTotals *d_tot = g->fb_totals.as<FbFileTotals>();
    PDG_CUDA(cudaMemcpyAsync(d_files, hf.data(), (n_files + 1) * sizeof(FbFile), cudaMemcpyHostToDevice, s));
    while (g->i_events.size() < 8) {
        cudaEvent_t ev;
        PDG_CUDA(cudaEventCreate(&ev));
        g->i_events.push_back(ev);
    }
    const uint64_t launches0 = g->launches;
    PDG_CUDA(cudaEventRecord(g->i_events[0], s));
    static const uint64_t group_bytes =
        getenv("PDG_FB_GROUP_BYTES") ? (uint64_t)atoll(getenv("PDG_FB_GROUP_BYTES")) : FB_GROUP_BYTES;
    static const bool no_pipeline = getenv("PDG_NO_PIPELINE") != nullptr;
    g->pre_valid = false;
    // k <= 11: msd3.cu on the whole code; k >= 17: msd3.cu on the 11-residue prefix (msdlong.cu finishes the buckets)
    if (!no_pipeline && (2 * g->k <= kMsd3MaxBits || g->k > 16) && g->n_res == 0 && g->c_begin.empty() &&
        g->x_src.empty() && g->x_nodes.empty() && tiles) {
        // device images are one group: nothing to overlap but the caller's own bookkeeping
        const int rc = ingest_pipelined(g, ptrs, n_bytes, n_files, h_table, cap_headers, infos, hf, dev, tiles,
                                        host_src ? group_bytes : ~0ULL, host_src);
        const bool launched = rc == PDG_OK && g->pre_valid;
        if (!launched) {   // else: i_events[5] was recorded before pass 3 was launched, and is complete
            PDG_CUDA(cudaEventRecord(g->i_events[5], s));
            PDG_CUDA(cudaEventSynchronize(g->i_events[5]));
        }
        g->ms_ingest[0] = g->ms_ingest[1] = g->ms_ingest[2] = 0.f;
        if (host_src) {
            cudaEventElapsedTime(&g->ms_ingest[0], g->i_events[0], g->i_events[5]);   // whole call
            if (launched) cudaEventElapsedTime(&g->ms_ingest[1], g->i_events[0], g->i_events[2]);   // copies only
        } else if (launched) {
            cudaEventElapsedTime(&g->ms_ingest[0], g->i_events[6], g->i_events[7]);   // pass A
            cudaEventElapsedTime(&g->ms_ingest[2], g->i_events[3], g->i_events[4]);   // pass C
        }
        uint64_t nb = 0;
        for (uint32_t f = 0; f < n_files; ++f) nb += n_bytes[f];
        g->ingest_bytes = nb;
        g->ingest_launches = g->launches - launches0;
        return rc;
    }
    // ---- pass A (host mode: per copy group, overlapping the next group's copy) ----
    if (host_src) {
        // the staging area may still be read by work queued on s; order the copies after it
        if (!g->fb_ev_ready) PDG_CUDA(cudaEventCreateWithFlags(&g->fb_ev_ready, cudaEventDisableTiming));
        PDG_CUDA(cudaEventRecord(g->fb_ev_ready, s));
        PDG_CUDA(cudaStreamWaitEvent(g->copy_stream, g->fb_ev_ready, 0));
        uint32_t f0 = 0;
        size_t ev_i = 0;
        while (f0 < n_files) {
            uint32_t f1 = f0;
            uint64_t bytes = 0;
            while (f1 < n_files && (f1 == f0 || bytes + n_bytes[f1] <= group_bytes)) {
                if (n_bytes[f1])
                    PDG_CUDA(cudaMemcpyAsync(const_cast<uint8_t *>(dev[f1]), ptrs[f1], n_bytes[f1],
                                             cudaMemcpyHostToDevice, g->copy_stream));
                bytes += n_bytes[f1];
                ++f1;
            }
            if (ev_i == g->fb_events.size()) {
                cudaEvent_t ev;
                PDG_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
                g->fb_events.push_back(ev);
            }
            PDG_CUDA(cudaEventRecord(g->fb_events[ev_i], g->copy_stream));
            PDG_CUDA(cudaStreamWaitEvent(s, g->fb_events[ev_i], 0));
            ++ev_i;
            const uint64_t tb = hf[f0].tile0, te = hf[f1].tile0;
            if (te > tb) {
                fb_summ_kernel<<<(unsigned)(te - tb), FB_THREADS, 0, s>>>(d_files, n_files, tb, d_summ);
                PDG_LAUNCHED(g);
            }
            f0 = f1;
        }
    } else if (tiles) {
        fb_summ_kernel<<<(unsigned)tiles, FB_THREADS, 0, s>>>(d_files, n_files, 0, d_summ);
        PDG_LAUNCHED(g);
    }
    PDG_CUDA(cudaEventRecord(g->i_events[1], s));
    fb_scan_kernel<<<n_files, FB_THREADS, 0, s>>>(d_files, 0, d_summ, d_prefix, d_tot);
    PDG_LAUNCHED(g);
    PDG_CUDA(cudaEventRecord(g->i_events[2], s));
    std::vector<FbFileTotals> tot(n_files);
    PDG_CUDA(cudaMemcpyAsync(tot.data(), d_tot, n_files * sizeof(FbFileTotals), cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));
    // ---- totals -> bases ----
    uint64_t n_rows = 0, n_res = 0;
    bool pure = true;
    for (uint32_t f = 0; f < n_files; ++f) {
        pdg_ingest_info o{};
        o.n_bytes = n_bytes[f];
        o.n_headers = tot[f].hdr;
        o.n_residues = tot[f].res;
        o.residue_base = g->n_res + n_res;
        o.appended = 1;
        infos[f] = o;
        hf[f].res_base = g->n_res + n_res;
        hf[f].hdr_base = n_rows;
        n_rows += tot[f].hdr;
        n_res += tot[f].res;
        pure = pure && tot[f].impure == 0;
    }
    hf[n_files].hdr_base = n_rows;
    if (n_rows > cap_headers) {
        set_error("pdg_ingest_fasta_batch: %llu headers, table capacity %llu", (unsigned long long)n_rows,
                  (unsigned long long)cap_headers);
        return PDG_E_CAPACITY;
    }
    if (!pure) return batch_fallback(g, dev, n_bytes, n_files, h_table, infos);
    // ---- pass C ----
    const uint64_t words = (g->n_res + n_res + 31) / 32 + 2;
    g->packed.reserve(words * 8, s, /*keep=*/true, /*zero=*/true);
    g->small.reserve((n_rows * 4 + 4) * sizeof(uint64_t), s);
    uint64_t *d_table = g->small.as<uint64_t>();
    PDG_CUDA(cudaMemcpyAsync(d_files, hf.data(), (n_files + 1) * sizeof(FbFile), cudaMemcpyHostToDevice, s));
    PDG_CUDA(cudaEventRecord(g->i_events[3], s));
    if (tiles) {
        fb_emit_kernel<<<(unsigned)tiles, FB_THREADS, 0, s>>>(d_files, n_files, 0, d_prefix,
                                                             g->packed.as<unsigned long long>(), d_table);
        PDG_LAUNCHED(g);
    }
    if (n_rows) {
        fb_hash_kernel<<<(unsigned)((n_rows + 127) / 128), 128, 0, s>>>(d_files, 0, n_files, d_tot, 0, n_rows, d_table);
        PDG_LAUNCHED(g);
    }
    PDG_CUDA(cudaEventRecord(g->i_events[4], s));
    if (n_rows)
        PDG_CUDA(cudaMemcpyAsync(h_table, d_table, n_rows * 4 * sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
    PDG_CUDA(cudaStreamSynchronize(s));   // hf / h_table are pageable host memory
    cudaEventElapsedTime(&g->ms_ingest[0], g->i_events[0], g->i_events[1]);
    cudaEventElapsedTime(&g->ms_ingest[1], g->i_events[1], g->i_events[2]);
    cudaEventElapsedTime(&g->ms_ingest[2], g->i_events[3], g->i_events[4]);
    uint64_t nb = 0;
    for (uint32_t f = 0; f < n_files; ++f) nb += n_bytes[f];
    g->ingest_bytes = nb;
    g->ingest_launches = g->launches - launches0;
    g->n_res += n_res;
    return PDG_OK;
}

}  // namespace pdg
