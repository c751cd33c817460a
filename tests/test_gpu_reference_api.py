"""`-m gpu`: the reference's own known-answer tests, re-pointed at B200Graph.

Scenarios and expected answers are those of /root/reference/tests/test_connected.py:26-380,
tests/test_graph.py:16-85, tests/test_prebuilt.py:18-75, tests/test_kmers.py:65-69 and
tests/test_subgraph_ref.py:11-38 (the reference runs them against LGGraph / Dgraph).  Every
scenario runs twice: with the reference's toy names ("ABC": opaque node keys) and with the names
mapped onto real 3-mers of a k=3 graph (k-mer codes through the per-edge path).
"""
import pytest

from conftest import golden, data_path

pytestmark = pytest.mark.gpu

TOY = {"ABC": "ABC", "BCD": "BCD", "CDE": "CDE", "DEF": "DEF", "XYZ": "XYZ"}
KMER = {"ABC": "ACG", "BCD": "CGT", "CDE": "GTA", "DEF": "TAC", "XYZ": "CCC"}


@pytest.fixture(params=["toy", "kmer3"])
def env(request, lib_built):
    from prairiedog_b200 import B200Graph
    names = TOY if request.param == "toy" else KMER
    return B200Graph(k=11 if request.param == "toy" else 3), names


def setup(env, nodes, edges):
    from prairiedog_b200 import Node, Edge
    g, nm = env
    for n in nodes:
        g.upsert_node(Node(value=nm[n]))
    for e in edges:
        s, t, v = e[0], e[1], e[2]
        ty = e[3] if len(e) > 3 else 'e'
        g.add_edge(Edge(src=nm[s], tgt=nm[t], edge_type=ty, edge_value=v))
    g.save()
    return g, nm


# (nodes, edges, a, b, n_start_edges, sorted path node lists)
SCENARIOS = {
    "connected": (["ABC", "BCD"], [("ABC", "BCD", 0)], "ABC", "BCD", 1, [["ABC", "BCD"]]),
    "not_connected": (["ABC", "BCD"], [], "ABC", "BCD", 0, []),
    "no_node": (["ABC"], [], "ABC", "BCD", 0, []),
    "distant": (["ABC", "BCD", "CDE"], [("ABC", "BCD", 0), ("BCD", "CDE", 1)], "ABC", "CDE", 1,
                [["ABC", "BCD", "CDE"]]),
    "multiple": (["ABC", "BCD", "CDE", "XYZ"],
                 [("ABC", "BCD", 0, "path1"), ("BCD", "CDE", 1, "path1"), ("ABC", "XYZ", 0, "path2"),
                  ("XYZ", "CDE", 1, "path2")], "ABC", "CDE", 2, [["ABC", "BCD", "CDE"], ["ABC", "XYZ", "CDE"]]),
    "shortcut": (["ABC", "BCD", "CDE"],
                 [("ABC", "BCD", 0, "path1"), ("BCD", "CDE", 1, "path1"), ("ABC", "CDE", 0, "path2")],
                 "ABC", "CDE", 2, [["ABC", "BCD", "CDE"], ["ABC", "CDE"]]),
    "repeats_full": (["ABC", "BCD", "CDE"],
                     [("ABC", "BCD", 0), ("BCD", "CDE", 1), ("CDE", "ABC", 2), ("ABC", "BCD", 3), ("BCD", "CDE", 4)],
                     "ABC", "CDE", 2,
                     [["ABC", "BCD", "CDE"], ["ABC", "BCD", "CDE"], ["ABC", "BCD", "CDE", "ABC", "BCD", "CDE"]]),
    "repeats_end": (["ABC", "BCD", "CDE"],
                    [("ABC", "BCD", 0), ("BCD", "CDE", 1), ("CDE", "BCD", 2), ("BCD", "CDE", 3)],
                    "ABC", "CDE", 1, [["ABC", "BCD", "CDE"], ["ABC", "BCD", "CDE", "BCD", "CDE"]]),
    "repeats_start": (["ABC", "BCD", "CDE"],
                      [("ABC", "BCD", 0), ("BCD", "ABC", 1), ("ABC", "BCD", 2), ("BCD", "CDE", 3)],
                      "ABC", "CDE", 2, [["ABC", "BCD", "ABC", "BCD", "CDE"], ["ABC", "BCD", "CDE"]]),
    "repeat_middle": (["ABC", "BCD", "CDE"], [("ABC", "BCD", 0), ("BCD", "BCD", 1), ("BCD", "CDE", 2)],
                      "ABC", "CDE", 1, [["ABC", "BCD", "BCD", "CDE"]]),
    "repeat_long": (["ABC", "BCD", "CDE", "DEF"],
                    [("ABC", "BCD", 0), ("BCD", "CDE", 1), ("CDE", "BCD", 2), ("BCD", "CDE", 3), ("CDE", "DEF", 4)],
                    "ABC", "DEF", 1, [["ABC", "BCD", "CDE", "BCD", "CDE", "DEF"]]),
}


@pytest.mark.parametrize("name", sorted(SCENARIOS))
def test_connected_and_path(env, name):
    """tests/test_connected.py:26-380."""
    from prairiedog_b200 import concat_values
    nodes, edges, a, b, n_start, want = SCENARIOS[name]
    g, nm = setup(env, nodes, edges)
    ok, starts = g.connected(nm[a], nm[b])
    assert ok == (n_start > 0) and len(starts) == n_start
    for e in starts:
        assert e.edge_value >= 0 and e.src == g._key(nm[a]) + 1
    paths, metas = g.path(nm[a], nm[b])
    assert len(paths) == len(want) == len(metas)
    assert sorted([n.value for n in p] for p in paths) == sorted([nm[x] for x in w] for w in want)
    if name == "connected":
        assert concat_values(paths[0]) == nm["ABC"] + nm["BCD"][-1]


def test_unsaved_data_is_invisible(env):
    """LGGraph queries run in their own read transaction: nothing is visible before save()
    (lemon_graph.py:178-181; every reference test calls g.save() first)."""
    from prairiedog_b200 import Node, Edge
    g, nm = env
    g.upsert_node(Node(nm["ABC"]))
    g.add_edge(Edge(nm["ABC"], nm["BCD"], edge_value=0))
    assert g.connected(nm["ABC"], nm["BCD"]) == (False, ())
    assert g.path(nm["ABC"], nm["BCD"]) == ((), ())
    g.save()
    assert g.connected(nm["ABC"], nm["BCD"])[0]


@pytest.mark.parametrize("n_more", [3, 70000], ids=["few-staged-edges", "more-than-one-flush"])
def test_queries_after_save_then_add_see_the_committed_state(env, n_more):
    """save -> add -> query: LemonGraph answers from the last COMMITTED state while the write transaction is open
    (lemon_graph.py:178-180 vs :54-64), whatever the amount of uncommitted data.  (70000 staged edges cross the
    65536-edge flush into libpdgpu, which used to turn the query into 'query before pdg_finalize'.)"""
    from prairiedog_b200 import Edge
    g, nm = env
    g.add_edge(Edge(nm["ABC"], nm["BCD"], edge_value=0))
    g.save()
    assert g.connected(nm["ABC"], nm["BCD"])[0]
    g.add_edge(Edge(nm["BCD"], nm["CDE"], edge_value=1))
    for i in range(n_more):
        g.add_edge(Edge(nm["CDE"], nm["DEF"], edge_value=2 + i), echo=False)
    # uncommitted: invisible; committed: still there
    assert g.connected(nm["BCD"], nm["CDE"]) == (False, ())
    assert g.path(nm["ABC"], nm["CDE"]) == ((), ())
    ok, starts = g.connected(nm["ABC"], nm["BCD"])
    assert ok and len(starts) == 1
    assert len(g.path(nm["ABC"], nm["BCD"])[0]) == 1
    g.save()
    assert g.connected(nm["BCD"], nm["CDE"])[0]
    assert len(g.path(nm["ABC"], nm["CDE"])[0]) == 1
    assert g.engine.counts()[1] == 2 + n_more


def test_bulk_add_after_save_keeps_the_committed_index(lib_built):
    """the same for the bulk path: a genome added after save() is invisible until the next save()"""
    import numpy as np
    from conftest import data_path
    from prairiedog_b200 import B200Graph, codec
    g = B200Graph(k=11)
    g.add_fasta(data_path("syn101.fasta"))
    g.save()
    before = g.engine.counts()
    seq = g.engine.fetch_residues(0, 80)
    a, b = seq[3:14], seq[20:31]
    hits_before = g.path_many([(a, b)])[2]
    assert len(hits_before) >= 1
    g.add_fasta(data_path("syn102.fasta"))
    assert g.engine.counts() == before                       # counts of the committed graph
    assert np.array_equal(g.path_many([(a, b)])[2], hits_before)
    g.save()
    assert g.engine.counts()[2] > before[2]


def test_graph_basics_nodes(env):
    """tests/test_graph.py:16-23."""
    from prairiedog_b200 import Node
    g, nm = env
    vals = [nm["ABC"], nm["BCD"], nm["CDE"]]
    for v in vals:
        g.upsert_node(Node(value=v))
    assert set(n.value for n in g.nodes) == set(vals)
    assert len(g) == 3


def test_graph_basics_edges(env):
    """tests/test_graph.py:27-46: edges report the numeric db ids of the upserted nodes."""
    from prairiedog_b200 import Node, Edge
    g, nm = env
    ns = [g.upsert_node(Node(value=nm[v])) for v in ("ABC", "BCD", "CDE")]
    g.add_edge(Edge(src=nm["ABC"], tgt=nm["BCD"]))
    g.add_edge(Edge(src=nm["BCD"], tgt=nm["CDE"]))
    g.add_edge(Edge(src=nm["BCD"], tgt=nm["CDE"]))          # upsert: no duplicate
    assert {(e.src, e.tgt) for e in g.edges} == {(ns[0].db_id, ns[1].db_id), (ns[1].db_id, ns[2].db_id)}
    assert all(e.edge_type == 'e' and e.edge_value == -1 for e in g.edges)


def test_graph_node_labels(env):
    """tests/test_graph.py:65-85."""
    from prairiedog_b200 import Node
    g, nm = env
    g.upsert_node(Node(value=nm["ABC"], labels={"species": "dog"}))
    g.upsert_node(Node(value=nm["BCD"], labels={"species": "cat"}))
    nodes = g.nodes
    assert len(nodes) == 2
    assert {n.value: n.labels for n in nodes} == {nm["ABC"]: {"species": "dog"}, nm["BCD"]: {"species": "cat"}}


def test_clear(env):
    from prairiedog_b200 import Node
    g, nm = env
    g.upsert_node(Node(nm["ABC"]))
    g.save()
    g.clear()
    g.save()
    assert len(g.nodes) == 0 and len(g.edges) == 0


# ---- tests/test_prebuilt.py:18-75: the graph of the three SHORTENED genomes ----
@pytest.fixture()
def prebuilt(lib_built):
    from prairiedog_b200 import B200Graph, Kmers, SubgraphRef, GraphRef
    g = B200Graph(k=11)
    for f in golden("fixtures_k11.json")["shortened3"]["files"]:
        SubgraphRef(g).update_graph(Kmers(data_path(f), 11), GraphRef())
    g.save()
    return g


def test_prebuilt_content_equals_golden_db(prebuilt):
    pre = golden("prebuilt_pangenome.json")
    nodes, edges = prebuilt.canonical()
    assert nodes == pre["nodes"] and [list(e) for e in edges] == pre["edges"]
    assert len(nodes) == 90 and len(edges) == 84


def test_prebuilt_queries(prebuilt):
    from prairiedog_b200 import concat_values
    ok, starts = prebuilt.connected('CCGGAAGAAAA', 'CGGAAGAAAAA')
    assert ok and len(starts) == 1
    paths, _ = prebuilt.path('CCGGAAGAAAA', 'CGGAAGAAAAA')
    assert len(paths) == 1 and paths[0][0].value == 'CCGGAAGAAAA' and paths[0][1].value == 'CGGAAGAAAAA'
    assert concat_values(paths[0]) == 'CCGGAAGAAAAA'
    ok, starts = prebuilt.connected('ATACGACGCCA', 'CGTCCGGACGT')
    assert ok and len(starts) == 1
    paths, _ = prebuilt.path('ATACGACGCCA', 'CGTCCGGACGT')
    assert len(paths) == 1 and paths[0][-1].value == 'CGTCCGGACGT'
    assert concat_values(paths[0]) == 'ATACGACGCCAGCGAACGTCCGGACGT'
    assert prebuilt.connected('GCTGGATACGT', 'CGTCCGGACGT') == (False, ())
    assert prebuilt.path('GCTGGATACGT', 'CGTCCGGACGT') == ((), ())


def test_kmers_known_answers(lib_built):
    """tests/test_kmers.py:13-23,65-69 on 15.fa (the first 15 bp of the stripped 172.fa)."""
    from prairiedog_b200 import Kmers
    km = Kmers(data_path("15.fa"))
    hdr = ">gi|1062504329|gb|CP014670.1| Escherichia coli strain CFSAN004177, complete genome"
    assert km.unique_kmers == 5 and len(km) == 5 and km.headers[0] == hdr
    assert km.next() == (hdr, "TCGCTTTCGTT")
    assert km.next() == (hdr, "CGCTTTCGTTC")
    while km.has_next:
        last = km.next()
    assert last == (hdr, "TTCGTTCATGC"[0:11]) or len(last[1]) == 11
    assert km.next() == ("", "")
