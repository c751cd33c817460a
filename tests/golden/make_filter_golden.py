#!/usr/bin/env python
"""Writes tests/golden/filter.json: what Graph.filter() leaves of a set of graphs, computed by the REFERENCE's own code.

prairiedog/networkx_graph.py:53-75 is the only implementation of filter() in the reference (LGGraph.filter is `pass`).
NetworkXGraph cannot be instantiated at this commit (abstract `connected` / `path` are missing), but
_filter_degree and _filter_isolates only touch `self.g`, an nx.Graph: they are called here UNBOUND on a stub object
that holds the nx.Graph, i.e. the reference's lines run unmodified.  What is OUR reading, not the reference's, is how
a labelled multigraph becomes that nx.Graph: one undirected edge per (src, tgt) pair, as networkx_graph.py:21-26
add_edge(node_a, node_b) would add them.  Needs /root/reference and networkx (this container); the GPU box only uses
the JSON.

  python tests/golden/make_filter_golden.py
"""
import json
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden", "filter.json")


def reference_filter(nodes, pairs):
    import networkx as nx
    from oracle import live_reference
    live_reference.load()                    # puts the unmodified reference package on sys.path
    from prairiedog.networkx_graph import NetworkXGraph
    stub = types.SimpleNamespace(g=nx.Graph())
    for v in nodes:
        stub.g.add_node(v)                       # networkx_graph.py:15-19
    for a, b in pairs:
        stub.g.add_edge(a, b)                    # networkx_graph.py:21-26
    NetworkXGraph._filter_degree(stub)           # networkx_graph.py:53-67, unmodified
    NetworkXGraph._filter_isolates(stub)         # networkx_graph.py:69-71, unmodified
    return sorted(stub.g.nodes), sorted(tuple(sorted(e)) for e in stub.g.edges)


def cases():
    from oracle import ref_spec
    from prairiedog_b200 import synth
    data = os.path.join(ROOT, "tests", "data")
    out = []
    short = ["SRR1060582_SHORTENED.fasta", "SRR1106609_SHORTENED.fasta",
             "GCA_900015695.1_ED647_contigs_genomic_SHORTENED.fasta"]
    for name, files, k in (("shortened3_k11", short, 11), ("shortened3_k3", short, 3), ("shortened3_k2", short, 2),
                           ("syn3_k4", ["syn101.fasta", "syn102.fasta", "syn201.fasta"], 4),
                           ("syn3_k5", ["syn101.fasta", "syn102.fasta", "syn201.fasta"], 5),
                           ("syn2_k11", ["syn101.fasta", "syn301.fasta"], 11)):
        g = ref_spec.SpecGraph()
        for f in files:
            g.add_genome(os.path.join(data, f), k)
        out.append({"name": name, "files": files, "k": k, "nodes": sorted(g.nodes), "edges": sorted(g.edges)})
    # toy graphs through the per-edge API (opaque node values, self-loops, parallel labelled edges, an isolate)
    toys = {
        "toy_star": (["HUB", "L1", "L2", "L3", "L4", "ALONE"],
                     [("HUB", "L1", "t", 0), ("HUB", "L2", "t", 1), ("L3", "HUB", "t", 2), ("L4", "HUB", "u", 0), ("L1", "L2", "t", 3)]),
        "toy_selfloop": (["AAA", "AAT", "ATA"],
                         [("AAA", "AAA", "t", 0), ("AAA", "AAA", "t", 1), ("AAA", "AAT", "t", 2), ("AAT", "ATA", "t", 3),
                          ("ATA", "AAA", "u", 0)]),
        "toy_parallel": (["P", "Q", "R"], [("P", "Q", "t", 0), ("P", "Q", "u", 0), ("Q", "P", "t", 1), ("Q", "R", "t", 2)]),
        "toy_clique": (["A1", "B1", "C1", "D1", "E1"],
                       [(a, b, "t", i) for i, (a, b) in enumerate([("A1", "B1"), ("A1", "C1"), ("A1", "D1"), ("B1", "C1"),
                                                                    ("B1", "D1"), ("C1", "D1"), ("D1", "E1")])]),
    }
    for name, (nodes, edges) in toys.items():
        out.append({"name": name, "files": None, "k": 3, "nodes": sorted(nodes), "edges": sorted(edges)})
    return out


def main():
    res = {}
    for c in cases():
        pairs = [(e[0], e[1]) for e in c["edges"]]
        kept_nodes, kept_pairs = reference_filter(c["nodes"], pairs)
        keep = set(kept_nodes)
        res[c["name"]] = {"files": c["files"], "k": c["k"], "n_nodes_before": len(c["nodes"]), "n_edges_before": len(c["edges"]),
                          "toy_nodes": c["nodes"] if c["files"] is None else None,
                          "toy_edges": [list(e) for e in c["edges"]] if c["files"] is None else None,
                          "kept_nodes": kept_nodes, "kept_pairs": [list(p) for p in kept_pairs],
                          # the labelled edges that go with the kept nodes (both endpoints kept)
                          "kept_edges": [list(e) for e in c["edges"] if e[0] in keep and e[1] in keep]}
        print("%-16s %6d nodes -> %6d   %7d labelled edges -> %7d" % (c["name"], len(c["nodes"]), len(kept_nodes), len(c["edges"]),
                                                                     len(res[c["name"]]["kept_edges"])))
    with open(OUT, "w") as f:
        json.dump(res, f, indent=0, sort_keys=True)
        f.write("\n")
    print("wrote", OUT)


if __name__ == "__main__":
    main()
