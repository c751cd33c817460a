"""`-m gpu`: B200Graph.filter() (pdg_filter_*) against tests/golden/filter.json -- what the reference's own
NetworkXGraph._filter_degree / _filter_isolates (prairiedog/networkx_graph.py:53-75), run unmodified on an nx.Graph by
tests/golden/make_filter_golden.py, leave of the same graphs -- and against the CPU restatement in oracle/ref_spec.py."""
import numpy as np
import pytest

from conftest import golden, data_path
from test_oracle_pinning import FILTER_CASES, spec_graph_of

pytestmark = pytest.mark.gpu


def b200_graph_of(case, general=False):
    from prairiedog_b200 import B200Graph, Node, Edge
    g = B200Graph(k=case["k"])
    if case["files"] is None:
        for v in case["toy_nodes"]:
            g.upsert_node(Node(value=v), echo=False)
        for s, t, ty, val in case["toy_edges"]:
            g.add_edge(Edge(src=s, tgt=t, edge_type=ty, edge_value=val), echo=False)
    else:
        for f in case["files"]:
            g.add_fasta(data_path(f))
    if general:                      # every contig expanded into the explicit part (CSR / CSC): the other edge source
        g._flush()
        g.engine.finalize(2)
        g._unsaved, g._published = False, True
    else:
        g.save()
    return g


@pytest.mark.parametrize("general", [False, True], ids=["index", "explicit"])
@pytest.mark.parametrize("name", FILTER_CASES)
def test_filter_equals_the_reference_code(lib_built, name, general):
    case = golden("filter.json")[name]
    g = b200_graph_of(case, general)
    nodes, edges = g.canonical()
    assert (len(nodes), len(edges)) == (case["n_nodes_before"], case["n_edges_before"])
    g.filter()
    nodes, edges = g.canonical()
    assert nodes == case["kept_nodes"]
    assert [list(e) for e in edges] == case["kept_edges"]
    assert len(g) == len(case["kept_nodes"])
    # the filtered graph is a committed graph like any other: queries answer from it
    if edges:
        s, t, ty, val = edges[0]
        ok, starts = g.connected(s, t)
        assert ok and any(e.edge_type == ty and e.edge_value == val for e in starts)
    # idempotence is NOT expected (the maximum degree changes), but a second pass must still agree with the oracle
    spec = spec_graph_of(case)
    spec.filter()
    spec.filter()
    g.filter()
    n2, e2 = g.canonical()
    assert (n2, e2) == spec.canonical()


def test_filter_of_an_empty_and_of_an_unsaved_graph(lib_built):
    from prairiedog_b200 import B200Graph, Edge
    g = B200Graph(k=11)
    g.filter()                                   # nothing saved: nothing to do, nothing raised
    assert len(g) == 0
    g.add_edge(Edge("A" * 11, "C" * 11, edge_type="t", edge_value=0))
    g.filter()                                   # commits first (like nodes / edges), then filters: one pair, both kept
    assert g.canonical() == (["A" * 11, "C" * 11], [("A" * 11, "C" * 11, "t", 0)])


def test_filter_at_scale_matches_the_oracle(lib_built):
    """2 synthetic genomes of 200 kbp at k = 11 (sparse: most nodes have degree 2) and at k = 6 (dense)"""
    from oracle import ref_spec
    from prairiedog_b200 import B200Graph, synth, codec
    for k in (11, 6):
        g = B200Graph(k=k)
        pairs = set()
        for name, buf, _ in synth.genomes(2, n_bases=200_000, n_contigs=8):
            g.add_fasta_bytes(name, buf)
        g.save()
        src, tgt, ty, val = g.engine.export_edges()
        keep = ref_spec.filter_keep(set(src.tolist()) | set(tgt.tolist()), zip(src.tolist(), tgt.tolist()))
        want = sorted((s, t, y, v) for s, t, y, v in zip(src.tolist(), tgt.tolist(), ty.tolist(), val.tolist())
                      if s in keep and t in keep)
        n_kept, fs, ft, fy, fv = g.engine.filter_edges()
        assert n_kept == len(keep)
        assert sorted(zip(fs.tolist(), ft.tolist(), fy.tolist(), fv.tolist())) == want
