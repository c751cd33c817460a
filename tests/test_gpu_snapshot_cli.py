"""`-m gpu`: persistent snapshot, exporters and the CLI driver against the
golden graphs of the reference (tests/golden)."""
import json
import os
import shutil

import numpy as np
import pytest

from conftest import golden, data_path, DATA

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _need_gpu(lib_built):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"


def _pair_on_one_contig(case, span=2):
    """(src, tgt, edge_type) with a path of `span` + 1 edges along one contig"""
    by = {(e[2], e[3]): e for e in case["edges"]}
    for (ty, v), e in sorted(by.items()):
        if (ty, v + span) in by:
            return e[0], by[(ty, v + span)][1], ty
    raise AssertionError("no contig with %d consecutive edges" % (span + 1))


def _build(case):
    from prairiedog_b200 import B200Graph
    g = B200Graph(k=case["k"])
    g.add_fasta_batch([(f, np.fromfile(data_path(f), dtype=np.uint8), None) for f in case["files"]])
    g.save()
    return g


@pytest.mark.parametrize("gold,name", [("fixtures_k11.json", "all4"), ("synthetic.json", "syn_k31"),
                                       ("synthetic.json", "syn_k5_repeats")])
def test_snapshot_round_trip_keeps_the_graph(tmp_path, gold, name):
    """snapshot(f) -> B200Graph.load(f): same canonical nodes / labelled edges as the reference, same
    query answers; appending to a loaded graph works (the store's tail stays zero)"""
    from prairiedog_b200 import B200Graph
    case = golden(gold)[name]
    g = _build(case)
    p = str(tmp_path / "pangenome.pdg")
    g.snapshot(p)
    h = B200Graph.load(p)
    nodes, edges = h.canonical()
    assert nodes == case["nodes"]
    assert [list(e) for e in edges] == case["edges"]
    assert h.engine.counts() == g.engine.counts()
    assert h.engine.checksum() == g.engine.checksum()
    src, tgt, ty = _pair_on_one_contig(case)
    assert {'edge_type': ty} in h.path(src, tgt)[1]
    assert sorted(m['edge_type'] for m in h.path(src, tgt)[1]) == sorted(m['edge_type'] for m in g.path(src, tgt)[1])
    assert h.connected(src, tgt)[0] == g.connected(src, tgt)[0]
    # a second snapshot of the loaded graph is byte-identical in its arrays
    q = str(tmp_path / "again.pdg")
    h.snapshot(q)
    from prairiedog_b200 import snapshot
    (m1, a1), (m2, a2) = snapshot.read(p), snapshot.read(q)
    assert all(np.array_equal(a1[n], a2[n]) for n in a1)
    assert m1["files"] == m2["files"] and m1["types"] == m2["types"]


def test_snapshot_with_explicit_part_and_labels(tmp_path):
    """per-edge API content (opaque node values, labels) survives the round trip"""
    from prairiedog_b200 import B200Graph, Edge, Node
    g = B200Graph(k=3)
    g.upsert_node(Node(value="ABC", labels={"colour": "red"}))
    g.add_edge(Edge(src="ABC", tgt="BCD", edge_type="path", edge_value=0, labels={"incr": 1}))
    g.add_edge(Edge(src="ATC", tgt="TCG", edge_type="g in f", edge_value=0))
    g.add_edge(Edge(src="TCG", tgt="CGA", edge_type="g in f", edge_value=1))
    g.set_graph_labels({"mic": 4})
    g.save()
    p = str(tmp_path / "x.pdg")
    g.snapshot(p)
    h = B200Graph.load(p)
    assert h.canonical() == g.canonical()
    assert h.get_labels("ABC") == g.get_labels("ABC")
    assert h.path("ATC", "CGA")[1] == ({'edge_type': 'g in f'},)
    assert h.connected("ABC", "BCD")[0]
    assert h._graph_labels == {"mic": 4}


def test_exporters(tmp_path):
    from prairiedog_b200 import snapshot
    case = golden("synthetic.json")["syn_k5_repeats"]
    g = _build(case)
    rdf, el = str(tmp_path / "g.rdf"), str(tmp_path / "g.edgelist")
    n = snapshot.write_rdf(g, rdf)
    assert n == len(case["edges"])
    lines = open(rdf).read().splitlines()
    assert len(lines) == len(case["nodes"]) + 4 * n
    assert lines[0] == '_:{v} <km> "{v}" .'.format(v=sorted(case["nodes"], key=lambda v: ["ATCG".index(c) for c in v])[0])
    want = set()
    for s, t, ty, v in case["edges"]:
        want.update(snapshot.rdf_lines(s, t, ty, v))
    assert set(lines[len(case["nodes"]):]) == want
    assert snapshot.write_edgelist(g, el) == n
    assert sorted(open(el).read().split("\n")[:-1]) == sorted("{} {}".format(e[0], e[1]) for e in case["edges"])


def test_cli_pangenome_query_and_batch(tmp_path):
    """Snakefile rules kmers + pangenome + done, then `prairiedog query` (cli.py:82-90)"""
    from click.testing import CliRunner
    from prairiedog_b200 import cli
    samples = tmp_path / "samples"
    samples.mkdir()
    case = golden("fixtures_k11.json")["all4"]
    for f in case["files"]:
        shutil.copy(data_path(f), samples / f)
    (samples / "notes.txt").write_text("not a genome")
    out = str(tmp_path / "outputs" / "pangenome.pdg")
    r = CliRunner().invoke(cli.cli, ["pangenome", str(samples), "--k", "11", "--out", out])
    assert r.exit_code == 0, r.output
    assert "{} nodes, {} edges".format(len(case["nodes"]), len(case["edges"])) in r.output
    assert os.path.exists(out)
    src, tgt, ty = _pair_on_one_contig(case)
    r = CliRunner().invoke(cli.cli, ["query", src, tgt, "--graph", out])
    assert r.exit_code == 0, r.output
    sample, contig = ty.split(" in ")[1], ty.split(" in ")[0]
    assert "Sample {} has:".format(sample) in r.output and "Contig {} with variant(s):".format(contig) in r.output
    pairs = tmp_path / "pairs.txt"
    pairs.write_text("# src dst\n{} {}\nAAAAAAAAAAA CCCCCCCCCCC\n".format(src, tgt))
    r = CliRunner().invoke(cli.cli, ["query-batch", str(pairs), "--graph", out])
    assert r.exit_code == 0, r.output
    rows = [json.loads(ln) for ln in r.output.splitlines() if ln.startswith("{")]
    assert len(rows) == 2 and rows[1]["hits"] == []
    assert any(h["edge_type"] == ty and h["string"].startswith(src) and h["string"].endswith(tgt)
               and len(h["string"]) == 11 + 3
               for h in rows[0]["hits"])
    r = CliRunner().invoke(cli.cli, ["export", "--graph", out, "--format", "edgelist", "--out", str(tmp_path / "e.txt")])
    assert r.exit_code == 0 and "{} edges".format(len(case["edges"])) in r.output
    r = CliRunner().invoke(cli.cli, ["query", src, tgt, "--backend", "dgraph", "--graph", out])
    assert r.exit_code != 0


@pytest.mark.parametrize("name", ["15_k11", "syn_k5"])
def test_update_graph_encode_true_matches_reference(name):
    """encode=True labels nodes by first-seen k-mer id (graph_ref.py:97-99);
    golden from the reference's LGGraph (tests/golden/make_encode_golden.py)"""
    from prairiedog_b200 import B200Graph, Kmers, SubgraphRef, GraphRef
    case = golden("encode.json")[name]
    g = B200Graph(k=case["k"])
    sg, gr = SubgraphRef(g), GraphRef()
    returned = []
    for f in case["files"]:
        km = Kmers(data_path(f), case["k"])
        gr.index_kmers(km)
        returned.append(sg.update_graph(km, gr, encode=True))
        sg.save()
    assert returned == case["returned"]
    assert gr.kmer_map == case["kmer_map"]
    nodes, edges = g.canonical()
    assert [repr(n) for n in nodes] == case["nodes"]
    assert [[repr(e[0]), repr(e[1]), e[2], e[3]] for e in edges] == case["edges"]
