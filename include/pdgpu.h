/*
 * pdgpu.h -- C ABI of libpdgpu.so, the B200 (sm_100a) pangenome-graph engine.
 *
 * This is the drop-in boundary for prairiedog's graph-construction hot path.  The reference has
 * no FFI of its own on this path except LemonGraph's cffi module; every entry point below names
 * the reference interface it replaces (paths relative to the reference checkout):
 *
 *   prairiedog/kmers.py:51-150          Kmers._load / next / _count_unique
 *   prairiedog/subgraph_ref.py:39-115   SubgraphRef.update_graph
 *   prairiedog/lemon_graph.py:99-334    LGGraph.upsert_node / add_edge / save / nodes / edges /
 *                                       connected / _find_path / path
 *   lemongraph/LemonGraph/cffi_stubs.py:100-236   graph_node_resolve / graph_edge_resolve /
 *                                       graph_txn_* / graph_nodes / graph_edges (native boundary)
 *
 * Conventions: plain C, no torch types.  Every function returns 0 on success and a negative
 * PDG_E_* code on failure; pdg_last_error() then describes the failure (thread local).  A graph
 * handle is used by one host thread at a time.  Pointers named d_* are device pointers (for
 * example torch.Tensor.data_ptr()), h_* are host pointers.  All work is enqueued on the stream
 * given to pdg_set_stream (default: the legacy default stream) and calls that return data to
 * the host synchronise that stream before returning.  There is no CPU implementation behind any
 * of these calls.
 *
 * Encoding: a k-mer over {A,T,C,G} is the integer sum(base[j] << 2*(k-1-j)) with A=0,T=1,C=2,G=3
 * (the enumeration order of kmers.py:17-24 possible_kmers), k <= 31.  Codes with bit 63 set are
 * opaque node keys interned by the host (node values that are not clean k-mers, for example the
 * three letter names of tests/test_connected.py).  edge_value is a signed 32-bit integer.
 *
 * Data model.  A graph holds two disjoint parts that together are the reference's node set and
 * labelled edge set:
 *   G  genome part: 2-bit packed residues + a contig table + the k-mer OCCURRENCE INDEX
 *      (positions of every k-mer occurrence, grouped by k-mer code, ascending position inside a
 *      group).  A labelled edge (src, tgt, edge_type, edge_value) of subgraph_ref.py:71-80 is a
 *      pure function of an occurrence position p inside contig c: src = kmer(p), tgt = kmer(p+1),
 *      edge_type = type[c], edge_value = p - begin[c]; so the same index serves as the by-source
 *      (out-edges: occurrences that are not last in their contig) and the by-target (in-edges:
 *      occurrences that are not first) adjacency of lemongraph.c:603-630.
 *   X  explicit part: edges / nodes given one at a time through the per-edge API, kept as
 *      deduplicated CSR + CSC over their own node table.
 * A contig whose type id is shared with another contig or with an explicit edge is expanded into
 * X at pdg_finalize so that set semantics (lemongraph.c:674-685) stay exact.
 */
#ifndef PDGPU_H
#define PDGPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PDG_ABI_VERSION 6

#define PDG_OK            0
#define PDG_E_INVALID    -1   /* bad argument / bad state */
#define PDG_E_CUDA       -2   /* a CUDA runtime call or kernel failed */
#define PDG_E_CAPACITY   -3   /* caller-provided buffer too small (required size reported) */
#define PDG_E_NOMEM      -4   /* device allocation failed */
#define PDG_E_COMM       -5   /* NCCL failure */

typedef struct pdg_graph pdg_graph;

/* result of pdg_ingest_fasta */
typedef struct pdg_ingest_info {
    uint64_t n_bytes;        /* bytes parsed */
    uint64_t n_headers;      /* lines starting with '>' (kmers.py:64-71) */
    uint64_t n_residues;     /* sequence characters after rstrip (kmers.py:63,72-73) */
    uint64_t n_dirty;        /* sequence bytes outside [ACGT] + interior blanks + bare CR */
    uint64_t residue_base;   /* index of the first residue of this file in the genome store */
    uint64_t appended;       /* 1 if residues were appended (n_dirty == 0), else 0 */
} pdg_ingest_info;

/* one path-query hit: a occurs as the source of edge (type_id, i), b as the target of edge
 * (type_id, j), i <= j, j - i <= max_len  (lemon_graph.py:275-334, graph.py:82-113).
 * aux: for a hit in the genome part, the index of the contig in pdg_add_contigs order among the
 * contigs kept in G (the path is residues [begin+i, begin+j+1+k) of that contig); for a hit in
 * the explicit part, PDG_HIT_EXPLICIT | node id (in the X node table) of the start edge's target. */
#define PDG_HIT_EXPLICIT 0x80000000u
typedef struct pdg_hit {
    uint32_t type_id;
    int32_t  i;
    int32_t  j;
    uint32_t aux;
} pdg_hit;

/* counters of the last pdg_finalize, for roofline accounting (bench.py) */
typedef struct pdg_build_stats {
    uint64_t n_residues, n_contigs, n_kmers, n_edge_records, n_nodes, n_edges;
    uint64_t node_path;      /* 0 = presence bitmap lookup, 1 = binary search over sorted codes */
    uint64_t n_demoted;      /* contigs expanded into the explicit part */
    uint64_t sort_passes;    /* partition passes run for the occurrence index */
    uint64_t sort_records;   /* records per pass */
    float ms_upload, ms_nodes, ms_sort, ms_rows, ms_explicit, ms_total;
    float ms_sort_kernel_avg;   /* CUDA-event average of one partition pass kernel */
    float pad;
    uint64_t sort_kernel_launches;
    uint64_t kernel_launches;   /* kernels launched on this handle since pdg_create / pdg_clear */
    /* CUDA-event times of the k <= 11 index build (msd3.cu): joint histogram, offset scans, and
     * the three partition passes; 0 when another path ran */
    float ms_hist, ms_scans, ms_pass[3], pad2;
    /* CUDA-event times of the last pdg_ingest_fasta_batch: pass A (summaries; in host mode it
     * includes waiting for the copies), pass B (scan), pass C (emit + header hashes) */
    float ms_ingest_a, ms_ingest_b, ms_ingest_c, pad3;
    uint64_t sort_groups;       /* groups of positions the k <= 11 index was staged in (1 unless pipelined) */
    uint64_t pipelined;         /* 1 if passes 1-2 already ran inside pdg_ingest_fasta_batch (host images) */
    uint64_t ingest_bytes;      /* bytes parsed by the last pdg_ingest_fasta_batch */
    uint64_t ingest_launches;
} pdg_build_stats;

const char *pdg_last_error(void);
int pdg_abi_version(void);
/* sizeof(pdg_hit), sizeof(pdg_ingest_info), sizeof(pdg_build_stats) as compiled into the library,
 * so that a binding can verify its struct mirrors (needs no device) */
int pdg_struct_sizes(uint64_t *hit, uint64_t *ingest_info, uint64_t *build_stats);

/* lifetime.  Replaces LGGraph.__init__ / clear (lemon_graph.py:27-42,130-131). */
pdg_graph *pdg_create(int device, int k);
void pdg_destroy(pdg_graph *g);
int pdg_set_stream(pdg_graph *g, void *cuda_stream);
int pdg_clear(pdg_graph *g);

/* K0+K1a: parse one FASTA file image that is resident on the device and append its residues,
 * packed to 2 bits, to the genome store.  Replaces Kmers._load (kmers.py:51-77).
 * h_header_table receives n_headers rows of {byte_begin, byte_end, residue_index}: the header
 * line is bytes [byte_begin, byte_end) of the file and residue_index residues precede it.
 * If n_headers > cap_headers the call fails with PDG_E_CAPACITY and info->n_headers holds the
 * required capacity.  If n_dirty != 0 nothing is appended (the host must route the file through
 * pdg_add_edges_host). */
int pdg_ingest_fasta(pdg_graph *g, const uint8_t *d_bytes, uint64_t n_bytes,
                     uint64_t *h_header_table, uint64_t cap_headers, pdg_ingest_info *info);

/* Batched pdg_ingest_fasta: n_files images in one call (three launches for the whole batch instead
 * of three per file), with the same results as calling pdg_ingest_fasta on every file in order --
 * same residue_base / n_residues / n_headers / n_dirty per file, same store contents.  ptrs[f] is
 * a device pointer, or with PDG_IN_HOST a host pointer (pinned memory recommended): the library
 * then stages the images on the device itself, copying on a second stream while the files that
 * already arrived are being parsed.  h_header_table receives the rows of all files back to back
 * (file f's rows start at the sum of infos[0..f).n_headers), FOUR uint64 per row: {byte_begin,
 * byte_end, residue_index, hash}; hash is a 64-bit hash of the right-stripped header line, or 0
 * when the host has to compare the bytes itself.  PDG_E_CAPACITY if the table is too small (every
 * infos[f].n_headers is then valid and nothing was appended).  Replaces the `for` loop over
 * genome files around Kmers._load (kmers.py:51-77; Snakefile rule kmers). */
#define PDG_IN_HOST 1
int pdg_ingest_fasta_batch(pdg_graph *g, const uint8_t *const *ptrs, const uint64_t *n_bytes,
                           uint32_t n_files, int flags, uint64_t *h_header_table,
                           uint64_t cap_headers, pdg_ingest_info *infos);

/* Declares contigs over residues already in the store: contig c covers residues
 * [h_begin[c], h_begin[c]+h_len[c]) and all its edges carry type id h_type_id[c].  Contigs with
 * h_len <= k yield nothing (kmers.py:110-117).  Replaces the per-contig bookkeeping of
 * SubgraphRef.update_graph (subgraph_ref.py:51-102). */
int pdg_add_contigs(pdg_graph *g, const uint64_t *h_begin, const uint64_t *h_len,
                    const uint32_t *h_type_id, uint64_t n_contigs);

/* Kmers.__len__ / unique_kmers (kmers.py:48-49,134-150) for a set of contigs in the store. */
int pdg_kmer_stats(pdg_graph *g, const uint64_t *h_begin, const uint64_t *h_len,
                   uint64_t n_contigs, uint64_t *n_kmers, uint64_t *n_unique);

/* k-mer codes of residues [begin, begin+len) (device output, n = len-k+1 codes); Kmers.next as
 * a bulk call.  pdg_fetch_residues decodes the same range back to ASCII (host output). */
int pdg_emit_codes(pdg_graph *g, uint64_t begin, uint64_t len, uint64_t *d_codes_out);
int pdg_fetch_residues(pdg_graph *g, uint64_t begin, uint64_t len, uint8_t *h_ascii);

/* Persistent snapshot.  Replaces the LMDB file LGGraph.save copies around
 * (lemon_graph.py:154-157, Snakefile:99-100): what is saved is the INPUT of pdg_finalize -- the
 * 2-bit store (0.25 B/bp), the declared contigs and the explicit edges / nodes -- because rebuilding
 * the index from it (tens of Gbp/s) is faster than reading an index back from disk.
 * pdg_snapshot_sizes: out = {n_residues, n_contigs, n_explicit_edges, n_explicit_nodes}.
 * pdg_snapshot_export copies them to host arrays of at least those sizes (h_words: at least
 * (n_residues + 31) / 32 words; residue i is bits 63-2*(i%32) .. of word i/32, A=0 T=1 C=2 G=3).
 * pdg_store_import replaces the store of an EMPTY graph (after pdg_create / pdg_clear); contigs,
 * edges and nodes are then declared with pdg_add_contigs / pdg_add_edges_host / pdg_add_nodes_host. */
int pdg_snapshot_sizes(pdg_graph *g, uint64_t *out4);
int pdg_snapshot_export(pdg_graph *g, uint64_t *h_words, uint64_t cap_words, uint64_t *h_begin,
                        uint64_t *h_len, uint32_t *h_type_id, uint64_t *h_src_code,
                        uint64_t *h_tgt_code, uint32_t *h_edge_type_id, int32_t *h_value,
                        uint64_t *h_node_code);
int pdg_store_import(pdg_graph *g, const uint64_t *h_words, uint64_t n_residues);

/* Per-edge compatibility path: LGGraph.add_edge / upsert_node (lemon_graph.py:99-128) and
 * graph_edge_resolve / graph_node_resolve (cffi_stubs.py).  Host arrays are copied. */
int pdg_add_edges_host(pdg_graph *g, const uint64_t *h_src_code, const uint64_t *h_tgt_code,
                       const uint32_t *h_type_id, const int32_t *h_value, uint64_t n);
int pdg_add_nodes_host(pdg_graph *g, const uint64_t *h_code, uint64_t n);

/* K2+K3: node dedup -> dense ids, occurrence index (G) and CSR + CSC (X).  Replaces the LMDB
 * upserts behind LGGraph.add_edge + LGGraph.save (lemon_graph.py:115-128,154-157;
 * lemongraph.c:661-685).  flags: bit0 = never use the bitmap lookup, bit1 = expand every contig
 * into the explicit part (general path, used by the parity tests). */
#define PDG_F_FORCE_SORT_NODES 1
#define PDG_F_FORCE_GENERAL    2
int pdg_finalize(pdg_graph *g, int flags);
/* Commit semantics: pdg_add_contigs / pdg_add_edges_host / pdg_add_nodes_host / pdg_ingest_fasta* after a
 * pdg_finalize do NOT invalidate its index -- queries, exports and pdg_counts keep answering from the last finalized
 * state until the next pdg_finalize, like a LemonGraph read transaction beside an open write transaction
 * (lemon_graph.py:178-180 vs :54-64).  Before the first pdg_finalize, and after pdg_clear, they fail / report an
 * empty graph. */
int pdg_counts(pdg_graph *g, uint64_t *n_nodes, uint64_t *n_edges, uint64_t *n_kmers);
int pdg_get_build_stats(pdg_graph *g, pdg_build_stats *out);

/* LGGraph.nodes / edges (lemon_graph.py:133-149).  Outputs are HOST arrays sized from
 * pdg_counts.  Nodes come in ascending code order (dense node id = index).  Edges: first the
 * genome part in index order (ascending (src code, position)), then the explicit part in CSR
 * order.  Any output pointer may be NULL.
 * In a multi-rank job (pdg_comm_init) pdg_finalize leaves the global node table distributed over the
 * ranks (every rank knows the global node COUNT); pdg_export_nodes, pdg_node_ids and pdg_checksum
 * gather it on their first call after pdg_finalize, so that call is COLLECTIVE: every rank makes it. */
int pdg_export_nodes(pdg_graph *g, uint64_t *h_codes);
int pdg_export_edges(pdg_graph *g, uint64_t *h_src_code, uint64_t *h_tgt_code,
                     uint32_t *h_type_id, int32_t *h_value);
/* dense node ids (rank in pdg_export_nodes order) of the given codes; UINT64_MAX if absent */
int pdg_node_ids(pdg_graph *g, const uint64_t *h_codes, uint64_t n, uint64_t *h_ids);
/* raw occurrence index of the genome part: h_row_codes[n_rows], h_row_ptr[n_rows+1],
 * h_positions[n_kmers] (sizes from pdg_index_size) */
int pdg_index_size(pdg_graph *g, uint64_t *n_rows, uint64_t *n_positions);
int pdg_export_index(pdg_graph *g, uint64_t *h_row_codes, uint32_t *h_row_ptr, uint32_t *h_positions);
/* order-independent 64-bit checksums of the canonical node set and edge set (sum of a mixed
 * hash over (code) / (src code, tgt code, type, value)), computed on the device from the index */
int pdg_checksum(pdg_graph *g, uint64_t *nodes_sum, uint64_t *edges_sum);

/* K4: batched queries.  h_a/h_b are node codes.  pdg_query_connected replaces LGGraph.connected
 * (lemon_graph.py:178-219): per query the start edges.  pdg_query_paths replaces
 * LGGraph.path (lemon_graph.py:275-334): hits per query with j - i <= max_len (max_len < 0: no
 * limit).  Results stay in graph-owned device memory until the next query call; pdg_query_fetch
 * copies them to the host (h_offsets has n_q + 1 entries).  pdg_query_walk materialises the node
 * codes of one hit's path (_find_path, lemon_graph.py:221-273): n = j - i + 2 codes; status 0 ok,
 * 1 chain broken, 2 ambiguous (the reference's `assert len(chains) == 1` would fail). */
int pdg_query_connected(pdg_graph *g, const uint64_t *h_a, const uint64_t *h_b, uint64_t n_q,
                        uint64_t *total_starts);
int pdg_query_paths(pdg_graph *g, const uint64_t *h_a, const uint64_t *h_b, uint64_t n_q,
                    int64_t max_len, uint64_t *total_hits);
/* pdg_query_fetch: either destination may also be DEVICE memory (a multi-rank caller exchanges the hits between the
 * ranks without a host round trip); the call returns after the copies have completed. */
int pdg_query_fetch(pdg_graph *g, uint64_t *h_offsets, pdg_hit *h_hits);
int pdg_query_walk(pdg_graph *g, uint64_t src_code, const pdg_hit *hit, uint64_t *h_node_codes,
                   int32_t *status);
/* A query whose hit count does not fit 32 bits fails the whole call with PDG_E_CAPACITY (nothing is written). */

/* The STRING of every hit of the last pdg_query_paths call in one go: Prairiedog.query spells a path with
 * concat_values (prairiedog/prairiedog.py:22-31, prairiedog/node.py:22-37) = the first node value followed by the
 * last character of every later one = residues [begin + i, begin + j + 1 + k) of the hit's contig, decoded from the
 * 2-bit store by one kernel (replaces one pdg_query_walk round trip per hit).
 * pdg_query_strings_size sizes the byte buffer; pdg_query_strings_fetch fills h_offsets[n_hits + 1] (byte offset of
 * hit h's string; starts are 16-byte aligned, so the LENGTH is j - i + 1 + k, not the difference of two offsets) and
 * h_bytes (ASCII, pinned memory recommended).  Hits of the explicit part (aux & PDG_HIT_EXPLICIT) have length 0:
 * their node values can be opaque host strings, spell them with pdg_query_walk. */
int pdg_query_strings_size(pdg_graph *g, uint64_t *total_bytes);
int pdg_query_strings_fetch(pdg_graph *g, uint64_t *h_offsets, uint8_t *h_bytes);

/* Graph.filter() (prairiedog/graph.py:62-63).  The one backend of the reference that implements it is NetworkXGraph
 * (prairiedog/networkx_graph.py:53-75; LGGraph.filter is `pass`): on the UNDIRECTED SIMPLE graph of the node pairs
 * -- one edge per pair whatever the labels, a self-loop counting twice in a degree -- remove the nodes with
 * degree <= int(0.9 * max degree), then the nodes left without a neighbour.  Three calls so that the threshold is
 * formed by the host exactly as the reference forms it (a Python float product truncated by int()):
 *   pdg_filter_degrees  projects every labelled edge of the finalized graph onto that simple graph (pair keys over
 *                       dense node ids, radix sort + dedup) and returns the maximum degree;
 *   pdg_filter_select   applies both removals for the given bottom_degrees and returns how many nodes survive and how
 *                       many labelled edges have both endpoints among them (the induced sub-multigraph);
 *   pdg_filter_fetch    copies those labelled edges to host arrays of n_kept_edges entries and frees the working set.
 * The graph itself is not modified: the caller rebuilds it from the fetched edges (B200Graph.filter does).
 * Single-rank graphs only. */
int pdg_filter_degrees(pdg_graph *g, uint32_t *max_degree);
int pdg_filter_select(pdg_graph *g, uint32_t bottom_degrees, uint64_t *n_kept_nodes, uint64_t *n_kept_edges);
int pdg_filter_fetch(pdg_graph *g, uint64_t *h_src_code, uint64_t *h_tgt_code, uint32_t *h_type_id,
                     int32_t *h_value);

/* K5: multi-GPU (one process per GPU).  The host exchanges the 128-byte id with torch.distributed.  Genomes shard
 * by sample and edges never move; inside pdg_finalize the ranks agree on the global node table: for k <= 15 one
 * ncclAllGather of the presence bitmaps + an OR, otherwise the 2k-bit key space is RANGE partitioned into `world`
 * equal slices (csrc/comm.cu:range_owner) and every rank sends each owner its slice of its sorted table with one
 * grouped ncclSend/ncclRecv exchange.  pdg_destroy destroys the communicator.  No reference counterpart
 * (SURVEY.md 2.2). */
int pdg_comm_unique_id(uint8_t *h_id128);
int pdg_comm_init(pdg_graph *g, const uint8_t *h_id128, int rank, int world);

/* K6: a checksum that allocates nothing.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  *d_sum = sum over d_codes[0, n) of mix64(code + golden ratio) modulo 2^64 -- for a graph's node table (the
 * codes pdg_export_nodes returns) exactly the nodes_sum of pdg_checksum.  d_codes and d_sum are DEVICE pointers on
 * `device`; two kernels are launched on `cuda_stream` (NULL = the default stream) and the call returns without
 * synchronising.  The per-CTA partial sums live in a 1 MiB file-scope __device__ array that every GPU loading the
 * library holds until exit (csrc/noalloc.cu), which bounds a call to PDG_STATIC_MAX_CODES codes: more returns
 * PDG_E_CAPACITY.  No cudaMalloc, cudaFree or allocating library routine is called.  All calls on a device share the
 * array: calls that overlap on the GPU (different streams) race and must be ordered by the caller. */
#define PDG_STATIC_MAX_CODES (1ull << 27)
int pdg_codes_checksum_static(int device, const uint64_t *d_codes, uint64_t n, uint64_t *d_sum, void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* PDGPU_H */
