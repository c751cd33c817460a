/*
 * lg_driver.c -- drives the REFERENCE's native graph engine (liblemongraph.so, compiled from
 * /root/reference/lemongraph by oracle/build_ref.py) through the same calls the reference's
 * Python makes per edge, as the CPU baseline of kind "reference".
 *
 * TEST / BENCH INFRASTRUCTURE -- never linked into the product.
 *
 * What it reproduces (paths relative to /root/reference):
 *   prairiedog/kmers.py:51-124            load + sliding window (well-formed FASTA only)
 *   prairiedog/subgraph_ref.py:51-105     per consecutive k-mer pair one add_edge, commit every
 *                                         333 k-mers
 *   prairiedog/lemon_graph.py:27-42       graph_open(nosync, noreadahead) + mapsize 200 GiB
 *   prairiedog/lemon_graph.py:115-121     add_edge = graph_node_resolve x2 ('km', kmer) +
 *                                         graph_edge_resolve(type = "{header} in {file}",
 *                                         value = str(i))
 * It leaves out the Python interpreter and cffi marshalling around those calls, so it is an
 * UPPER bound on the reference's single-process throughput.
 *
 * usage: lg_driver <db path> <k> <max_bases> <fasta>...      (one JSON line on stdout)
 */
#include <fcntl.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "lemongraph.h"

int db_set_mapsize(db_t db, size_t mapsize);

static double now(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

int main(int argc, char **argv) {
    if (argc < 5) {
        fprintf(stderr, "usage: %s db k max_bases fasta...\n", argv[0]);
        return 2;
    }
    const char *db = argv[1];
    const int k = atoi(argv[2]);
    const long long max_bases = atoll(argv[3]);
    graph_t g = graph_open(db, O_RDWR | O_CREAT, 0760, DB_NOSYNC | DB_NORDAHEAD);
    if (!g) { perror("graph_open"); return 1; }
    db_set_mapsize((db_t)g, ((size_t)20 << 30) * 10);
    graph_txn_t txn = graph_txn_begin(g, NULL, 0);
    if (!txn) { perror("graph_txn_begin"); return 1; }
    long long bases = 0, edges = 0, kmers = 0, commits = 0;
    const double t0 = now();
    char *seq = NULL, *hdr = NULL, *type = NULL;
    size_t seq_cap = 0;
    for (int a = 4; a < argc && bases < max_bases; ++a) {
        FILE *f = fopen(argv[a], "r");
        if (!f) { perror(argv[a]); return 1; }
        const char *base = strrchr(argv[a], '/');
        base = base ? base + 1 : argv[a];
        char *line = NULL;
        size_t cap = 0, seq_len = 0;
        ssize_t n;
        int eof = 0;
        free(hdr); hdr = NULL;
        while (!eof && bases < max_bases) {
            n = getline(&line, &cap, f);
            if (n < 0) { eof = 1; n = 0; }
            while (n > 0 && (line[n - 1] == '\n' || line[n - 1] == '\r' || line[n - 1] == ' ' || line[n - 1] == '\t')) n--;
            if (eof || (n > 0 && line[0] == '>')) {
                /* flush the contig: kmers.py:96-124 + subgraph_ref.py:51-105 */
                if (hdr && seq_len > (size_t)k) {
                    const size_t tl = strlen(hdr) + 4 + strlen(base);
                    type = (char *)realloc(type, tl + 1);
                    sprintf(type, "%s in %s", hdr, base);
                    node_t src = graph_node_resolve(txn, (void *)"km", 2, seq, k);
                    kmers++;
                    for (size_t p = 1; p + k <= seq_len; ++p) {
                        node_t tgt = graph_node_resolve(txn, (void *)"km", 2, seq + p, k);
                        char val[24];
                        const int vl = sprintf(val, "%zu", p - 1);
                        edge_t e = graph_edge_resolve(txn, src, tgt, type, tl, val, vl);
                        free(e);
                        free(src);
                        src = tgt;
                        edges++;
                        kmers++;
                        if (kmers % 333 == 0) {          /* subgraph_ref.py:103-105 */
                            graph_txn_commit(txn);
                            txn = graph_txn_begin(g, NULL, 0);
                            commits++;
                        }
                    }
                    free(src);
                    bases += (long long)seq_len;
                }
                seq_len = 0;
                if (!eof) {
                    free(hdr);
                    hdr = strndup(line, (size_t)n);
                }
            } else if (n > 0) {
                if (seq_len + (size_t)n + 1 > seq_cap) {
                    seq_cap = (seq_len + (size_t)n + 1) * 2;
                    seq = (char *)realloc(seq, seq_cap);
                }
                memcpy(seq + seq_len, line, (size_t)n);
                seq_len += (size_t)n;
            }
        }
        free(line);
        fclose(f);
    }
    const size_t nn = graph_nodes_count(txn, 0), ne = graph_edges_count(txn, 0);
    graph_txn_commit(txn);
    const double dt = now() - t0;
    graph_close(g);
    printf("{\"bases\": %lld, \"kmers\": %lld, \"edges\": %lld, \"commits\": %lld, \"seconds\": %.6f, "
           "\"lg_nodes\": %zu, \"lg_edges\": %zu, \"bp_per_s\": %.1f}\n",
           bases, kmers, edges, commits, dt, nn, ne, bases / dt);
    return 0;
}
