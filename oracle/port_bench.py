"""Times oracle/pdg_oracle.c on a block of the synthetic genomes (bench.py's cpu_baseline.port leg).

TEST / BENCH INFRASTRUCTURE: the product never imports this.  One process = one thread = one graph.
  python -m oracle.port_bench FIRST COUNT K BASES CONTIGS RELATED   -> one JSON line {bases, seconds, ...}
The genomes are generated first; the timed region is FASTA parse + sliding window + sort + dedup (orc_add_fasta_mem
for every genome, then orc_build), i.e. what prairiedog/kmers.py:51-124 + subgraph_ref.py:51-102 compute."""
import json
import sys
import time


def main():
    first, count, k, bases, contigs, related = (int(x) for x in sys.argv[1:7])
    from oracle.c_oracle import COracle, build
    from prairiedog_b200 import synth
    build()
    bufs, n_bases = [], 0
    for _, buf, lens in synth.genomes(count, n_bases=bases, n_contigs=contigs, first=first, related=bool(related)):
        bufs.append(buf)
        n_bases += int(lens.sum())
    t0 = time.time()
    o = COracle(k)
    for b in bufs:
        o.add_fasta_bytes(b)
    o.build()
    dt = time.time() - t0
    n_nodes, n_edges, n_kmers = o.counts()
    print(json.dumps({"bases": n_bases, "seconds": dt, "n_nodes": n_nodes, "n_edges": n_edges, "n_kmers": n_kmers, "k": k}))


if __name__ == "__main__":
    main()
