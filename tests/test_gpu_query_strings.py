"""`-m gpu`: pdg_query_strings (one kernel spells every hit of a batch) against three independent spellings:
concat_values over the node chain of pdg_query_walk (the per-hit path it replaces, prairiedog/node.py:22-37), the
contig slice read back with pdg_fetch_residues, and the reference's own strings for the golden query pairs."""
import bisect

import numpy as np
import pytest

from conftest import golden, data_path

pytestmark = pytest.mark.gpu


def _pairs_from(g, rng, n, max_d):
    """random (a, b) k-mer pairs d apart on the genome store (many land inside one contig: positives)"""
    n_res = int(g.engine.stats()["n_residues"])
    seq = g.engine.fetch_residues(0, n_res)
    out = []
    for _ in range(n):
        p = int(rng.integers(0, n_res - g.k - max_d - 1))
        d = int(rng.integers(1, max_d + 1))
        out.append((seq[p:p + g.k], seq[p + d:p + d + g.k]))
    return out


@pytest.mark.parametrize("k", [5, 11, 16, 17, 31])
def test_strings_equal_walk_and_contig_slice(lib_built, k):
    from prairiedog_b200 import B200Graph
    g = B200Graph(k=k)
    for f in ("syn101.fasta", "syn102.fasta", "syn201.fasta"):
        g.add_fasta(data_path(f))
    g.save()
    rng = np.random.default_rng(k)
    pairs = _pairs_from(g, rng, 60, 70)
    a, off, hits = g.path_many(pairs, limit=90)
    assert hits.size > 0
    soff, sbytes = g.engine.query_strings()
    assert soff.size == hits.size + 1 and (soff % 16 == 0).all()
    raw = sbytes.tobytes()
    qof = np.repeat(np.arange(len(pairs)), np.diff(off.astype(np.int64)))
    lens_seen = set()
    for x, h in enumerate(hits):
        ln = int(h["j"]) - int(h["i"]) + 1 + k
        lens_seen.add(ln % 16)
        s = raw[int(soff[x]):int(soff[x]) + ln].decode("ascii")
        # (1) the per-hit path it replaces
        codes, status = g.engine.walk(int(a[qof[x]]), h)
        assert status == 0
        vals = [g._value(c) for c in codes.tolist()]
        assert s == vals[0] + "".join(v[-1] for v in vals[1:])
        # (2) the contig slice itself
        f = g._files[bisect.bisect_right(g._file_bases, int(h["type_id"])) - 1]      # type id -> file -> contig
        begin = f.residue_base + int(f.seq_begin[int(h["type_id"]) - f.type_base])
        assert s == g.engine.fetch_residues(begin + int(h["i"]), ln)
        # padding up to the next 16-byte start is zero, never a neighbour's characters
        pad = raw[int(soff[x]) + ln:int(soff[x + 1])]
        assert pad == b"\0" * len(pad)
        assert s.startswith(pairs[qof[x]][0]) and s.endswith(pairs[qof[x]][1])
    assert len(lens_seen) > 4          # lengths of many residues mod 16 were exercised


def test_query_many_strings_match_the_reference_golden(lib_built):
    """the three query pairs of the reference's tests/test_prebuilt.py with the strings the unmodified reference
    returned (tests/golden/prebuilt_pangenome.json)"""
    from prairiedog_b200 import B200Graph
    case = golden("fixtures_k11.json")["shortened3"]
    g = B200Graph(k=case["k"])
    for f in case["files"]:
        g.add_fasta(data_path(f))
    g.save()
    qs = [q for q in case["queries"] if "raises" not in q]
    res = g.query_many([(q["a"], q["b"]) for q in qs])
    assert sum(len(q["hits"]) for q in qs) > 0
    for q, row in zip(qs, res):
        # golden hits: the (string, edge_type) pairs Prairiedog.query of the unmodified reference produced
        assert sorted((d["string"], d["edge_type"]) for d in row) == sorted(tuple(h) for h in q["hits"])
        for d in row:
            assert d["string"].startswith(q["a"]) and d["string"].endswith(q["b"])
            assert len(d["string"]) == d["j"] - d["i"] + 1 + case["k"]


def test_strings_of_an_empty_batch_and_of_connected(lib_built):
    from prairiedog_b200 import B200Graph, _ffi
    g = B200Graph(k=11)
    g.add_fasta(data_path("syn101.fasta"))
    g.save()
    a, off, hits = g.path_many([("A" * 11, "C" * 11)])
    if hits.size == 0:
        soff, sbytes = g.engine.query_strings()
        assert soff.tolist() == [0] and sbytes.size == 0
    g.connected_many([("A" * 11, "C" * 11)])
    with pytest.raises(_ffi.PdgError):          # start edges are not paths: no strings
        g.engine.query_strings()
