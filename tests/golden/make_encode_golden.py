#!/usr/bin/env python
"""Golden vectors for SubgraphRef.update_graph(..., encode=True) (subgraph_ref.py:54,65;
graph_ref.py:97-99): node values are first-seen-order k-mer ids instead of k-mer strings.
Runs the UNMODIFIED reference (LGGraph backend) in the build container; writes encode.json."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
DATA = os.path.join(ROOT, "tests", "data")


def main():
    from oracle import live_reference as L
    ns = L.load()
    out = {}
    for name, files, k in (("15_k11", ["15.fa"], 11), ("syn_k5", ["syn201.fasta", "syn202.fasta"], 5)):
        g = L.new_lggraph()
        sg, gr = ns.SubgraphRef(g), ns.GraphRef()
        cwd = os.getcwd()
        os.chdir(ns.scratch)
        returned = []
        try:
            for f in files:
                km = ns.Kmers(os.path.join(DATA, f), k)
                gr.index_kmers(km)
                returned.append(sg.update_graph(km, gr, encode=True))
                sg.save(None)
        finally:
            os.chdir(cwd)
        nodes, edges = L.canonical(g)
        out[name] = {"files": files, "k": k, "returned": returned, "kmer_map": gr.kmer_map,
                     "nodes": [repr(n) for n in nodes], "edges": [[repr(e[0]), repr(e[1]), e[2], e[3]] for e in edges]}
        print(name, len(nodes), "nodes", len(edges), "edges", nodes[:3])
    json.dump(out, open(os.path.join(HERE, "encode.json"), "w"), indent=0)


if __name__ == "__main__":
    main()
