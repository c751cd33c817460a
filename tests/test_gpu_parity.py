"""`-m gpu`: the CUDA path (through the C ABI) against the oracle and the golden vectors of the
unmodified reference.  Bar: bit-exact (integer / byte work)."""
import os

import numpy as np
import pytest

from conftest import golden, data_path

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _need_gpu(lib_built):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"


def build_b200(files, k, flags=0):
    from prairiedog_b200 import B200Graph
    g = B200Graph(k=k)
    returned = [g.add_fasta(data_path(f)) for f in files]
    g._flush()
    g.engine.finalize(flags)
    g._unsaved, g._published = False, True
    return g, returned


def c_oracle(files, k):
    from oracle.c_oracle import COracle
    o = COracle(k)
    for f in files:
        o.add_fasta(data_path(f))
    o.build()
    return o


CASES = [("fixtures_k11.json", n) for n in ["15.fa", "SRR1060582_SHORTENED.fasta", "SRR1106609_SHORTENED.fasta",
                                            "GCA_900015695.1_ED647_contigs_genomic_SHORTENED.fasta",
                                            "shortened3", "all4"]] + \
        [("synthetic.json", n) for n in ["syn_k11", "syn_k5_repeats", "syn_k16", "syn_k17", "syn_k31", "syn_k3"]]


@pytest.mark.parametrize("gold,name", CASES)
@pytest.mark.parametrize("flags", [0, 1, 2], ids=["index", "nobitmap", "general"])
def test_graph_matches_reference(gold, name, flags):
    """node set, labelled edge set, update_graph return value (configs 1 and 2 of BASELINE.json)."""
    case = golden(gold)[name]
    g, returned = build_b200(case["files"], case["k"], flags)
    assert returned == [s["returned"] for s in case["stats"]]
    nodes, edges = g.canonical()
    assert nodes == case["nodes"]
    assert [list(e) for e in edges] == case["edges"]
    n_nodes, n_edges, n_kmers = g.engine.counts()
    assert (n_nodes, n_edges) == (len(case["nodes"]), len(case["edges"]))
    assert n_kmers == sum(s["len"] for s in case["stats"])
    if flags != 2:
        o = c_oracle(case["files"], case["k"])
        assert g.engine.checksum() == o.checksum()
        for got, want in zip(g.engine.export_index(), o.index()):
            assert np.array_equal(got, want)
        for got, want in zip(g.engine.export_edges(), o.edges()):
            assert np.array_equal(got, want)


@pytest.mark.parametrize("gold,name", CASES)
@pytest.mark.parametrize("flags", [0, 2], ids=["index", "general"])
def test_queries_match_reference(gold, name, flags):
    from prairiedog_b200 import Prairiedog, concat_values
    case = golden(gold)[name]
    g, _ = build_b200(case["files"], case["k"], flags)
    for q in case["queries"]:
        if "raises" in q:       # the reference crashed there
            continue
        ok, starts = g.connected(q["a"], q["b"])
        assert ok == q["connected"] and len(starts) == q["n_start"]
        paths, metas = g.path(q["a"], q["b"])
        got = sorted([concat_values(p), m["edge_type"]] for p, m in zip(paths, metas))
        assert got == q["hits"]
        for p in paths:
            assert p[0].value == q["a"] and p[-1].value == q["b"]
        hits = Prairiedog(g).query(q["a"], q["b"])
        assert sorted([h["string"], h["edge_type"]] for h in hits) == q["hits"]


@pytest.mark.parametrize("gold,name", CASES)
def test_kmers_stats(gold, name):
    from prairiedog_b200 import Kmers
    case = golden(gold)[name]
    for f, s in zip(case["files"], case["stats"]):
        km = Kmers(data_path(f), case["k"])
        assert (len(km), km.unique_kmers) == (s["len"], s["unique"])
        assert str(km) == f


def test_per_edge_api_equals_bulk():
    """The reference's own loop (one add_edge per k-mer pair) through the per-edge API gives the
    same graph as the bulk path, also when both are mixed in one graph (set semantics)."""
    from prairiedog_b200 import B200Graph, Kmers, SubgraphRef, GraphRef, Edge
    from oracle import ref_spec
    case = golden("synthetic.json")["syn_k11"]
    g = B200Graph(k=11)
    for f in case["files"]:
        _, edges, _ = ref_spec.genome_edges(data_path(f), 11)
        for s, t, ty, v in edges:
            g.add_edge(Edge(s, t, ty, v), echo=False)
    g.save()
    nodes, edges = g.canonical()
    assert nodes == case["nodes"] and [list(e) for e in edges] == case["edges"]
    # bulk-load the same files on top: every edge already exists, nothing may change
    for f in case["files"]:
        SubgraphRef(g).update_graph(Kmers(data_path(f), 11), GraphRef())
    g.save()
    assert g.engine.stats()["n_demoted"] > 0
    nodes2, edges2 = g.canonical()
    assert nodes2 == nodes and edges2 == edges
    for q in case["queries"]:
        assert len(g.path(q["a"], q["b"])[0]) == len(q["hits"])


def test_same_file_twice_is_idempotent():
    case = golden("fixtures_k11.json")["shortened3"]
    g, _ = build_b200(case["files"] + case["files"][:1], 11)
    nodes, edges = g.canonical()
    assert nodes == case["nodes"] and [list(e) for e in edges] == case["edges"]


@pytest.mark.parametrize("name", sorted(golden("corner_cases.json")))
def test_corner_cases(name):
    """FASTA corner cases of SURVEY.md A.1 through Kmers + SubgraphRef.update_graph (k=4)."""
    from prairiedog_b200 import B200Graph, Kmers, SubgraphRef, GraphRef
    rec = golden("corner_cases.json")[name]
    path = data_path(os.path.join("corner", name + ".fasta"))
    if "raises" in rec:
        with pytest.raises(IndexError):
            Kmers(path, 4)
        return
    km = Kmers(path, 4)
    assert km.headers == rec["headers"] and km.sequences == rec["sequences"]
    assert (len(km), km.unique_kmers) == (rec["len"], rec["unique"])
    em = []
    while km.has_next:
        em.append(list(km.next()))
    km.reset()
    assert em == rec["emitted"]
    g = B200Graph(k=4)
    assert SubgraphRef(g).update_graph(km, GraphRef()) == rec["returned"]
    g.save()
    nodes, edges = g.canonical()
    assert nodes == rec["nodes"] and [list(e) for e in edges] == rec["edges"]
