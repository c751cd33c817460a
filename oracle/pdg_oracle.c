/*
 * pdg_oracle.c -- CPU restatement (plain C) of prairiedog's graph-construction hot path.
 *
 * TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library; the product (prairiedog_b200, libpdgpu.so) never
 * does.  Parity status: PINNED -- tests/test_oracle_pinning.py checks it, together with
 * oracle/ref_spec.py, against the golden vectors produced by the unmodified reference
 * (tests/golden/, made by tests/golden/make_golden.py) and the known answers in the reference's
 * own tests.
 *
 * What is restated (paths relative to /root/reference):
 *   orc_add_fasta   prairiedog/kmers.py:51-77     Kmers._load: rstrip each line, '>' lines are
 *                                                 headers, other lines are concatenated
 *   orc_build       prairiedog/kmers.py:96-124    sliding window over every contig with len > k
 *                   prairiedog/subgraph_ref.py:62-102  edge (kmer_i -> kmer_i+1, type, value=i)
 *                   lemongraph/lib/lemongraph.c:661-685  nodes unique on value, edges unique on
 *                                                 (type, value, src, tgt)
 *   orc_query       prairiedog/lemon_graph.py:178-334 + prairiedog/graph.py:82-113
 * Only well-formed input is handled (upper-case ACGT, every record non-empty, no two adjacent
 * contigs with len <= k, last contig long): anything else returns an error and the tests use
 * oracle/ref_spec.py, which follows the reference through its corner cases.
 *
 * Build: gcc -O2 -shared -fPIC oracle/pdg_oracle.c -o oracle/_build/libpdgoracle.so
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    uint64_t begin, len;
    uint32_t type;
} orc_contig;

typedef struct {
    uint64_t code;
    uint32_t pos;
} orc_occ;

typedef struct orc_graph {
    int k;
    uint8_t *res;            /* residues as 2-bit digits, one per byte: A=0 T=1 C=2 G=3 */
    uint64_t n_res, cap_res;
    orc_contig *contigs;
    uint64_t n_contigs, cap_contigs;
    /* built */
    orc_occ *occ;            /* sorted by (code, pos) */
    uint64_t n_occ;
    uint64_t *row_code;
    uint64_t *row_ptr;
    uint64_t n_rows, n_edges;
} orc_graph;

static int digit_of(int c) {
    switch (c) {
        case 'A': return 0;
        case 'T': return 1;
        case 'C': return 2;
        case 'G': return 3;
        default: return -1;
    }
}

orc_graph *orc_new(int k) {
    if (k < 1 || k > 31) return NULL;
    orc_graph *g = (orc_graph *)calloc(1, sizeof(orc_graph));
    g->k = k;
    return g;
}

void orc_free(orc_graph *g) {
    if (!g) return;
    free(g->res); free(g->contigs); free(g->occ); free(g->row_code); free(g->row_ptr);
    free(g);
}

static void push_res(orc_graph *g, uint8_t d) {
    if (g->n_res == g->cap_res) {
        g->cap_res = g->cap_res ? g->cap_res * 2 : (1u << 20);
        g->res = (uint8_t *)realloc(g->res, g->cap_res);
    }
    g->res[g->n_res++] = d;
}

static void push_contig(orc_graph *g, uint64_t begin, uint64_t len, uint32_t type) {
    if (g->n_contigs == g->cap_contigs) {
        g->cap_contigs = g->cap_contigs ? g->cap_contigs * 2 : 1024;
        g->contigs = (orc_contig *)realloc(g->contigs, g->cap_contigs * sizeof(orc_contig));
    }
    g->contigs[g->n_contigs].begin = begin;
    g->contigs[g->n_contigs].len = len;
    g->contigs[g->n_contigs].type = type;
    g->n_contigs++;
}

/* kmers.py:61-75.  Contig c of the file gets type id type_base + c.  Returns the number of
 * headers, or -1 file error, -2 alphabet, -3 irregular layout. */
int64_t orc_add_fasta_mem(orc_graph *g, const uint8_t *buf, uint64_t n, uint32_t type_base) {
    const uint64_t res0 = g->n_res, contigs0 = g->n_contigs;
    int64_t n_headers = 0;
    uint64_t seq_begin = g->n_res;
    int have_header = 0;
    uint64_t i = 0;
    while (i < n) {
        uint64_t e = i;
        while (e < n && buf[e] != '\n') e++;
        uint64_t r = e;                                  /* rstrip (kmers.py:63) */
        while (r > i && (buf[r - 1] == '\r' || buf[r - 1] == ' ' || buf[r - 1] == '\t' ||
                         buf[r - 1] == '\v' || buf[r - 1] == '\f')) r--;
        if (r > i && buf[i] == '>') {
            if (g->n_res != seq_begin) {                 /* flush the previous sequence */
                if (!have_header) goto irregular;        /* text before the first header */
                push_contig(g, seq_begin, g->n_res - seq_begin, type_base + (uint32_t)(n_headers - 1));
                seq_begin = g->n_res;
            } else if (have_header) {
                goto irregular;                          /* header without a sequence */
            }
            have_header = 1;
            n_headers++;
        } else {
            for (uint64_t j = i; j < r; ++j) {
                const int d = digit_of(buf[j]);
                if (d < 0) { g->n_res = res0; g->n_contigs = contigs0; return -2; }
                push_res(g, (uint8_t)d);
            }
        }
        i = e + 1;
    }
    if (!have_header || g->n_res == seq_begin) goto irregular;
    push_contig(g, seq_begin, g->n_res - seq_begin, type_base + (uint32_t)(n_headers - 1));
    /* the iterator of kmers.py:96-124 simply skips short contigs only in this arrangement */
    for (uint64_t c = contigs0; c < g->n_contigs; ++c) {
        const int s = g->contigs[c].len <= (uint64_t)g->k;
        if (s && (c + 1 == g->n_contigs || g->contigs[c + 1].len <= (uint64_t)g->k)) goto irregular;
    }
    return n_headers;
irregular:
    g->n_res = res0;
    g->n_contigs = contigs0;
    return -3;
}

int64_t orc_add_fasta(orc_graph *g, const char *path, uint32_t type_base) {
    FILE *f = fopen(path, "rb");
    if (!f) return -1;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    uint8_t *buf = (uint8_t *)malloc(n > 0 ? (size_t)n : 1);
    if (n > 0 && fread(buf, 1, (size_t)n, f) != (size_t)n) { fclose(f); free(buf); return -1; }
    fclose(f);
    const int64_t r = orc_add_fasta_mem(g, buf, (uint64_t)n, type_base);
    free(buf);
    return r;
}

/* stable LSD radix sort of (code, pos) records on the code, 11 bits per pass */
static void sort_occ(orc_occ *a, uint64_t n, int key_bits) {
    orc_occ *b = (orc_occ *)malloc((n ? n : 1) * sizeof(orc_occ));
    uint64_t *cnt = (uint64_t *)malloc(2049 * sizeof(uint64_t));
    for (int shift = 0; shift < key_bits; shift += 11) {
        memset(cnt, 0, 2049 * sizeof(uint64_t));
        for (uint64_t i = 0; i < n; ++i) cnt[((a[i].code >> shift) & 2047) + 1]++;
        for (int d = 0; d < 2048; ++d) cnt[d + 1] += cnt[d];
        for (uint64_t i = 0; i < n; ++i) b[cnt[(a[i].code >> shift) & 2047]++] = a[i];
        memcpy(a, b, n * sizeof(orc_occ));
    }
    free(b);
    free(cnt);
}

/* kmers.py:96-124 + subgraph_ref.py:62-102 + lemongraph.c:661-685 */
int orc_build(orc_graph *g) {
    const int k = g->k;
    const uint64_t mask = (k == 32) ? ~0ULL : ((1ULL << (2 * k)) - 1ULL);
    free(g->occ); free(g->row_code); free(g->row_ptr);
    uint64_t n = 0;
    g->n_edges = 0;
    for (uint64_t c = 0; c < g->n_contigs; ++c)
        if (g->contigs[c].len > (uint64_t)k) {
            n += g->contigs[c].len - k + 1;
            g->n_edges += g->contigs[c].len - k;
        }
    g->n_occ = n;
    g->occ = (orc_occ *)malloc((n ? n : 1) * sizeof(orc_occ));
    uint64_t o = 0;
    for (uint64_t c = 0; c < g->n_contigs; ++c) {
        const orc_contig *ct = &g->contigs[c];
        if (ct->len <= (uint64_t)k) continue;            /* kmers.py:110-117 */
        uint64_t code = 0;
        for (uint64_t p = 0; p < ct->len; ++p) {
            code = ((code << 2) | g->res[ct->begin + p]) & mask;
            if (p + 1 >= (uint64_t)k) {
                g->occ[o].code = code;
                g->occ[o].pos = (uint32_t)(ct->begin + p + 1 - k);
                o++;
            }
        }
    }
    sort_occ(g->occ, n, 2 * k);
    uint64_t rows = 0;
    for (uint64_t i = 0; i < n; ++i)
        if (i == 0 || g->occ[i].code != g->occ[i - 1].code) rows++;
    g->n_rows = rows;
    g->row_code = (uint64_t *)malloc((rows ? rows : 1) * sizeof(uint64_t));
    g->row_ptr = (uint64_t *)malloc((rows + 1) * sizeof(uint64_t));
    rows = 0;
    for (uint64_t i = 0; i < n; ++i)
        if (i == 0 || g->occ[i].code != g->occ[i - 1].code) {
            g->row_code[rows] = g->occ[i].code;
            g->row_ptr[rows] = i;
            rows++;
        }
    g->row_ptr[rows] = n;
    return 0;
}

void orc_counts(const orc_graph *g, uint64_t *n_nodes, uint64_t *n_edges, uint64_t *n_kmers) {
    *n_nodes = g->n_rows;
    *n_edges = g->n_edges;
    *n_kmers = g->n_occ;
}

void orc_index(const orc_graph *g, uint64_t *row_code, uint32_t *row_ptr, uint32_t *pos) {
    for (uint64_t r = 0; r < g->n_rows; ++r) row_code[r] = g->row_code[r];
    for (uint64_t r = 0; r <= g->n_rows; ++r) row_ptr[r] = (uint32_t)g->row_ptr[r];
    for (uint64_t i = 0; i < g->n_occ; ++i) pos[i] = g->occ[i].pos;
}

static const orc_contig *contig_of(const orc_graph *g, uint64_t p) {
    uint64_t lo = 0, hi = g->n_contigs;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (g->contigs[mid].begin <= p) lo = mid + 1; else hi = mid;
    }
    return &g->contigs[lo ? lo - 1 : 0];
}

static uint64_t code_at(const orc_graph *g, uint64_t p) {
    uint64_t c = 0;
    for (int j = 0; j < g->k; ++j) c = (c << 2) | g->res[p + j];
    return c;
}

/* edges in (src code, position) order; any output may be NULL; returns the count */
uint64_t orc_edges(const orc_graph *g, uint64_t *src, uint64_t *tgt, uint32_t *type, int32_t *val) {
    uint64_t e = 0;
    for (uint64_t i = 0; i < g->n_occ; ++i) {
        const uint64_t p = g->occ[i].pos;
        const orc_contig *ct = contig_of(g, p);
        if (p + g->k >= ct->begin + ct->len) continue;  /* last k-mer of its contig: no out-edge */
        if (src) src[e] = g->occ[i].code;
        if (tgt) tgt[e] = code_at(g, p + 1);
        if (type) type[e] = ct->type;
        if (val) val[e] = (int32_t)(p - ct->begin);
        e++;
    }
    return e;
}

static uint64_t mix64(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}

/* the order-independent sums pdg_checksum reports (include/pdgpu.h) */
void orc_checksum(const orc_graph *g, uint64_t *nodes_sum, uint64_t *edges_sum) {
    uint64_t ns = 0, es = 0;
    for (uint64_t r = 0; r < g->n_rows; ++r) ns += mix64(g->row_code[r] + 0x9e3779b97f4a7c15ULL);
    for (uint64_t i = 0; i < g->n_occ; ++i) {
        const uint64_t p = g->occ[i].pos;
        const orc_contig *ct = contig_of(g, p);
        if (p + g->k >= ct->begin + ct->len) continue;
        uint64_t h = mix64(g->occ[i].code + 0x9e3779b97f4a7c15ULL);
        h = mix64(h ^ code_at(g, p + 1));
        h = mix64(h ^ (((uint64_t)ct->type << 32) | (uint32_t)(int32_t)(p - ct->begin)));
        es += h;
    }
    *nodes_sum = ns;
    *edges_sum = es;
}

static int64_t find_row(const orc_graph *g, uint64_t code) {
    uint64_t lo = 0, hi = g->n_rows;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (g->row_code[mid] < code) lo = mid + 1; else hi = mid;
    }
    return (lo < g->n_rows && g->row_code[lo] == code) ? (int64_t)lo : -1;
}

/* lemon_graph.py:178-334 by direct enumeration: every out-edge (T,i) of a against every in-edge
 * (T,j) of b.  connected != 0: one hit per start edge that has some j >= i.  hits: rows of
 * {type, i, j}; returns the number of hits (writes at most cap). */
uint64_t orc_query(const orc_graph *g, uint64_t a, uint64_t b, int64_t max_len, int connected,
                   int64_t *hits, uint64_t cap) {
    const int64_t ra = find_row(g, a), rb = find_row(g, b);
    if (ra < 0 || rb < 0) return 0;
    uint64_t n = 0;
    for (uint64_t x = g->row_ptr[ra]; x < g->row_ptr[ra + 1]; ++x) {
        const uint64_t p = g->occ[x].pos;
        const orc_contig *ca = contig_of(g, p);
        if (p + g->k >= ca->begin + ca->len) continue;   /* not an out-edge */
        const int64_t i = (int64_t)(p - ca->begin);
        int any = 0;
        for (uint64_t y = g->row_ptr[rb]; y < g->row_ptr[rb + 1]; ++y) {
            const uint64_t q = g->occ[y].pos;
            const orc_contig *cb = contig_of(g, q);
            if (q == cb->begin) continue;               /* first k-mer: not an in-edge */
            if (cb->type != ca->type) continue;         /* graph.py:103 same edge_type */
            const int64_t j = (int64_t)(q - cb->begin) - 1;
            if (i > j) continue;                        /* graph.py:106 */
            if (connected) { any = 1; continue; }
            if (max_len >= 0 && j - i > max_len) continue;   /* lemon_graph.py:229-236 */
            if (n < cap) { hits[3 * n] = ca->type; hits[3 * n + 1] = i; hits[3 * n + 2] = j; }
            n++;
        }
        if (connected && any) {
            if (n < cap) { hits[3 * n] = ca->type; hits[3 * n + 1] = i; hits[3 * n + 2] = i; }
            n++;
        }
    }
    return n;
}

/* ---- streaming forms for inputs too large to sort in one piece (tests/golden/make_bench_checksums.py) ----
 * They restate the SAME definitions as orc_build + orc_checksum without materialising the occurrence list;
 * tests/test_oracle_pinning.py checks them against orc_build / orc_checksum on the fixtures.
 * Valid for what orc_add_fasta accepts with every contig carrying its own type id (no (type, value) pair can
 * repeat, so the labelled edge SET of lemongraph.c:661-685 is one edge per non-final k-mer occurrence). */

/* n_off edge sums at once: sums[o] uses type id (contig type + type_off[o]) -- one pass over the residues */
void orc_stream_sums(const orc_graph *g, const uint32_t *type_off, int n_off, uint64_t *n_kmers, uint64_t *n_edges,
                     uint64_t *sums) {
    const int k = g->k;
    const uint64_t mask = (1ULL << (2 * k)) - 1ULL;
    uint64_t nk = 0, ne = 0;
    for (int o = 0; o < n_off; ++o) sums[o] = 0;
    for (uint64_t c = 0; c < g->n_contigs; ++c) {
        const orc_contig *ct = &g->contigs[c];
        if (ct->len <= (uint64_t)k) continue;            /* kmers.py:110-117 */
        uint64_t code = 0, prev = 0;
        for (uint64_t p = 0; p < ct->len; ++p) {
            code = ((code << 2) | g->res[ct->begin + p]) & mask;
            if (p + 1 < (uint64_t)k) continue;
            nk++;
            if (p + 1 > (uint64_t)k) {                   /* edge prev -> code, value = start of prev */
                const uint64_t val = p - (uint64_t)k;    /* (p + 1 - k) - 1 */
                const uint64_t h2 = mix64(mix64(prev + 0x9e3779b97f4a7c15ULL) ^ code);
                for (int o = 0; o < n_off; ++o)
                    sums[o] += mix64(h2 ^ (((uint64_t)(ct->type + type_off[o]) << 32) | (uint32_t)(int32_t)val));
                ne++;
            }
            prev = code;
        }
    }
    *n_kmers = nk;
    *n_edges = ne;
}

/* presence bits of every emitted k-mer code, OR-ed into bitmap (4^k bits, 64 per word); k <= 15 */
int orc_mark_codes(const orc_graph *g, uint64_t *bitmap) {
    const int k = g->k;
    if (k > 15) return -1;
    const uint64_t mask = (1ULL << (2 * k)) - 1ULL;
    for (uint64_t c = 0; c < g->n_contigs; ++c) {
        const orc_contig *ct = &g->contigs[c];
        if (ct->len <= (uint64_t)k) continue;
        uint64_t code = 0;
        for (uint64_t p = 0; p < ct->len; ++p) {
            code = ((code << 2) | g->res[ct->begin + p]) & mask;
            if (p + 1 >= (uint64_t)k) bitmap[code >> 6] |= 1ULL << (code & 63);
        }
    }
    return 0;
}

/* count and node checksum of the codes whose bit is set */
void orc_bitmap_sum(const uint64_t *bitmap, uint64_t n_words, uint64_t *n_nodes, uint64_t *nodes_sum) {
    uint64_t n = 0, s = 0;
    for (uint64_t w = 0; w < n_words; ++w)
        for (uint64_t m = bitmap[w]; m; m &= m - 1) {
            const uint64_t code = (w << 6) | (uint64_t)__builtin_ctzll(m);
            s += mix64(code + 0x9e3779b97f4a7c15ULL);
            n++;
        }
    *n_nodes = n;
    *nodes_sum = s;
}

/* every emitted k-mer code of the loaded contigs, in emission order; returns the count (out may be NULL) */
uint64_t orc_emit_codes(const orc_graph *g, uint64_t *out) {
    const int k = g->k;
    const uint64_t mask = (1ULL << (2 * k)) - 1ULL;
    uint64_t n = 0;
    for (uint64_t c = 0; c < g->n_contigs; ++c) {
        const orc_contig *ct = &g->contigs[c];
        if (ct->len <= (uint64_t)k) continue;
        uint64_t code = 0;
        for (uint64_t p = 0; p < ct->len; ++p) {
            code = ((code << 2) | g->res[ct->begin + p]) & mask;
            if (p + 1 >= (uint64_t)k) { if (out) out[n] = code; n++; }
        }
    }
    return n;
}

/* in place: LSD radix sort of n keys on key_bits bits (tmp: n words), then adjacent dedup; returns the count */
uint64_t orc_sort_unique(uint64_t *a, uint64_t *tmp, uint64_t n, int key_bits) {
    uint64_t *cnt = (uint64_t *)malloc(2049 * sizeof(uint64_t));
    uint64_t *src = a, *dst = tmp;
    for (int shift = 0; shift < key_bits; shift += 11) {
        memset(cnt, 0, 2049 * sizeof(uint64_t));
        for (uint64_t i = 0; i < n; ++i) cnt[((src[i] >> shift) & 2047) + 1]++;
        for (int d = 0; d < 2048; ++d) cnt[d + 1] += cnt[d];
        for (uint64_t i = 0; i < n; ++i) dst[cnt[(src[i] >> shift) & 2047]++] = src[i];
        uint64_t *t = src; src = dst; dst = t;
    }
    free(cnt);
    uint64_t m = 0;
    for (uint64_t i = 0; i < n; ++i)
        if (i == 0 || src[i] != src[i - 1]) a[m++] = src[i];   /* a == src or a == dst: writes trail reads */
    return m;
}

/* union of two sorted duplicate-free arrays into out (room for na + nb); returns the count */
uint64_t orc_merge_unique(const uint64_t *a, uint64_t na, const uint64_t *b, uint64_t nb, uint64_t *out) {
    uint64_t i = 0, j = 0, m = 0;
    while (i < na && j < nb) {
        if (a[i] < b[j]) out[m++] = a[i++];
        else if (b[j] < a[i]) out[m++] = b[j++];
        else { out[m++] = a[i++]; j++; }
    }
    while (i < na) out[m++] = a[i++];
    while (j < nb) out[m++] = b[j++];
    return m;
}

uint64_t orc_codes_sum(const uint64_t *codes, uint64_t n) {
    uint64_t s = 0;
    for (uint64_t i = 0; i < n; ++i) s += mix64(codes[i] + 0x9e3779b97f4a7c15ULL);
    return s;
}
