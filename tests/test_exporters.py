"""CPU: exporters and the hit printer against golden vectors written by the UNMODIFIED reference
(tests/golden/make_rdf_golden.py -> rdf.json), and the snapshot container format."""
import json
import os

import numpy as np

from conftest import golden


def test_rdf_lines_equal_the_reference_bulk_writer():
    """DgraphBulk.add_edge + save (prairiedog/dgraph.py:576-613)"""
    from prairiedog_b200 import snapshot
    for e in golden("rdf.json")["edges"]:
        assert list(snapshot.rdf_lines(e["src"], e["tgt"], e["edge_type"], e["edge_value"])) == e["lines"]


def test_pretty_hits_equal_the_reference():
    """prairiedog/pretty_hits.py:13-51: sample -> contig -> set of variants"""
    from prairiedog_b200.prairiedog import PrettyHits
    h = golden("rdf.json")["hits"]
    ph = PrettyHits(h["list"])
    assert {s: {c: sorted(v) for c, v in m.items()} for s, m in ph.sample_map.items()} == h["sample_map"]
    text = str(ph)
    assert "Sample g2.fasta has:" in text and "Contig >contig_7 with variant(s):" in text


def test_snapshot_container_round_trip(tmp_path):
    """header + aligned arrays; a reader rejects foreign files"""
    from prairiedog_b200 import snapshot

    class _Eng:
        nodes = True

        def snapshot_export(self):
            return {"words": np.arange(5, dtype=np.uint64), "c_begin": np.array([0, 70], np.uint64),
                    "c_len": np.array([60, 80], np.uint64), "c_type": np.array([0, 1], np.uint32),
                    "x_src": np.zeros(0, np.uint64), "x_tgt": np.zeros(0, np.uint64), "x_type": np.zeros(0, np.uint32),
                    "x_val": np.array([], np.int32), "x_nodes": np.array([1 << 63] if self.nodes else [], np.uint64),
                    "n_residues": 150}

    class _G:
        k = 11
        engine = _Eng()
        _files, _types, _opaque, _explicit_suffixes = [], {3: "a in b"}, ["ABC"], {"b"}
        _node_meta, _edge_labels, _graph_labels, _next_type = {5: ("km", {"x": 1})}, {(1, 2, 3, 4): {"incr": 2}}, {"mic": 1}, 4

        def _flush(self):
            pass
    p = str(tmp_path / "g.pdg")
    snapshot.save(_G(), p)
    meta, arr = snapshot.read(p)
    assert meta["k"] == 11 and meta["n_residues"] == 150 and meta["types"] == [[3, "a in b"]]
    assert arr["words"].tolist() == [0, 1, 2, 3, 4] and arr["c_len"].tolist() == [60, 80]
    assert arr["x_nodes"].tolist() == [1 << 63] and arr["x_val"].size == 0
    assert all(d["offset"] % 64 == 0 for d in meta["arrays"].values())
    _Eng.nodes = False                      # trailing arrays empty: nothing follows the last payload byte
    snapshot.save(_G(), p)
    meta, arr = snapshot.read(p)
    assert arr["x_nodes"].size == 0 and arr["c_type"].tolist() == [0, 1]
    bad = str(tmp_path / "bad.pdg")
    open(bad, "wb").write(b"not a snapshot at all")
    try:
        snapshot.read(bad)
        assert False
    except ValueError:
        pass


def test_graph_ref_mic_labels_and_first_seen_ids(tmp_path):
    """graph_ref.py:20-103 / gref.py:9-19: MIC values get ids from 1 in first-seen order, k-mers and
    files from 0; one-hot rows"""
    from prairiedog_b200.graph_ref import GraphRef

    class _Km:
        def __init__(self, p, u):
            self.filepath, self.unique_kmers = p, u
    csv = tmp_path / "mic.csv"
    csv.write_text("run,AMP,CIP\nSRR1,8.0,0.5\nSRR2,>=32,0.5\n")
    gr = GraphRef(str(csv))
    a, b = _Km("samples/SRR1.fasta", 10), _Km("x/SRR2.fa", 30)
    gr._record_graph_label(a)
    gr._record_graph_label(b)
    assert gr.mic_map == {"8.0": 1, "0.5": 2, ">=32": 3} or list(gr.mic_map.values()) == [1, 2, 3]
    assert gr.get_graph_label(b, "CIP") == gr.get_graph_label(a, "CIP")
    assert gr.get_graph_label(b, "AMP") == 3
    gr.index_kmers(a)
    gr.index_kmers(b)
    assert gr.max_num_nodes == 30
    assert [gr.node_label(x) for x in ("ATC", "TCG", "ATC")] == [0, 1, 0]
    assert gr.edge_label(a) == 0 and gr.edge_label(b) == 1 and gr.edge_label(a) == 0
    assert GraphRef._one_hot(2, 4).tolist() == [0, 0, 1, 0]
    assert GraphRef(None).MIC_DF is None


def test_cli_sample_files_follow_the_snakefile(tmp_path):
    """Snakefile:22-24: every *.fna / *.fasta / *.fa of the samples directory, nothing else; sorted so that
    type ids are reproducible"""
    from prairiedog_b200 import cli
    for name in ("b.fasta", "a.fa", "c.fna", "notes.txt", "d.fasta.gz", "public_mic_class_dataframe.csv"):
        (tmp_path / name).write_text(">x\nACGT\n")
    assert [p.split("/")[-1] for p in cli.sample_files(str(tmp_path))] == ["a.fa", "b.fasta", "c.fna"]
