#!/usr/bin/env python
"""Golden vectors for the RDF exporter and PrettyHits: the N-Quads the UNMODIFIED reference's
DgraphBulk.add_edge writes (prairiedog/dgraph.py:576-613) and str(PrettyHits(...))
(prairiedog/pretty_hits.py) for a few edges / hits.  Run in the build container
(needs /root/reference); writes tests/golden/rdf.json."""
import json
import os
import sys
import tempfile
import types

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def main():
    from oracle import live_reference
    ns = live_reference.load()
    # pydgraph / grpc are only needed by the live Dgraph client, not by the bulk writer
    stub = types.ModuleType("pydgraph")
    for name in ("DgraphClientStub", "DgraphClient", "Txn"):
        setattr(stub, name, type(name, (), {}))
    sys.modules.setdefault("pydgraph", stub)
    cwd = os.getcwd()
    os.chdir(ns.scratch)
    try:
        from prairiedog.dgraph import DgraphBulk
        from prairiedog.pretty_hits import PrettyHits
    finally:
        os.chdir(cwd)
    edges = [("ATCGATCGATC", "TCGATCGATCA", ">contig_1 len=9 in g1.fasta", 0),
             ("TCGATCGATCA", "CGATCGATCAG", ">contig_1 len=9 in g1.fasta", 1),
             ("GGGGGGGGGGG", "GGGGGGGGGGG", ">c2 in other.fa", 17),
             ("ABC", "BCD", "path", -1)]
    out = {"edges": [], "hits": None}
    for s, t, ty, v in edges:
        db = DgraphBulk()
        db.add_edge(ns.Edge(src=s, tgt=t, edge_type=ty, edge_value=v))
        f = os.path.join(tempfile.mkdtemp(prefix="pdg_rdf_"), "x.rdf")
        db.save(f)
        lines = [ln.strip() for ln in open(f).read().splitlines() if ln.strip()]
        out["edges"].append({"src": s, "tgt": t, "edge_type": ty, "edge_value": v, "lines": lines})
    hits = [{"string": "ATCGATCGATCAG", "edge_type": ">contig_1 len=9 in g1.fasta"},
            {"string": "ATCGATCGATCAG", "edge_type": ">contig_7 in g2.fasta"},
            {"string": "ATCGTTCGATCAG", "edge_type": ">contig_7 in g2.fasta"}]
    ph = PrettyHits(hits)
    out["hits"] = {"list": hits, "sample_map": {s: {c: sorted(v) for c, v in m.items()} for s, m in ph.sample_map.items()}}
    json.dump(out, open(os.path.join(HERE, "rdf.json"), "w"), indent=1)
    print("wrote rdf.json:", len(out["edges"]), "edges")


if __name__ == "__main__":
    main()
