"""`not gpu`: the C-ABI library loads and exports every symbol include/pdgpu.h declares; host-side
logic (codec, FASTA layout, API value classes) behaves like the reference's."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT, golden, data_path


def header_symbols():
    hdr = open(os.path.join(ROOT, "include", "pdgpu.h")).read()
    return sorted(set(re.findall(r"\b(pdg_[a-z_0-9]+)\s*\(", hdr)))


def test_library_exports_header_symbols(lib_built):
    from prairiedog_b200 import _ffi
    lib = ctypes.CDLL(lib_built)
    names = header_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(_ffi.SIGNATURES) == names
    lib.pdg_abi_version.restype = ctypes.c_int
    assert lib.pdg_abi_version() == 6


def test_struct_layouts(lib_built):
    from prairiedog_b200 import _ffi
    from prairiedog_b200.engine import HIT_DTYPE
    assert ctypes.sizeof(_ffi.Hit) == 16 == HIT_DTYPE.itemsize
    assert ctypes.sizeof(_ffi.IngestInfo) == 48
    assert ctypes.sizeof(_ffi.BuildStats) == 200
    import numpy as np
    assert np.dtype(_ffi.INGEST_INFO_DTYPE).itemsize == 48
    lib = _ffi.load()
    a, b, c = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64()
    assert lib.pdg_struct_sizes(ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)) == 0
    assert (a.value, b.value, c.value) == (ctypes.sizeof(_ffi.Hit), ctypes.sizeof(_ffi.IngestInfo),
                                           ctypes.sizeof(_ffi.BuildStats))


def test_no_device_fails_loudly(lib_built):
    """No CPU fallback: without a CUDA device pdg_create returns NULL and Engine raises."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from prairiedog_b200 import _ffi
    from prairiedog_b200.engine import Engine
    with pytest.raises(_ffi.PdgError):
        Engine(11, 0)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "prairiedog_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.replace("no oracle", ""), f


def test_codec_matches_reference_enumeration():
    import itertools
    from prairiedog_b200 import codec
    for k in (1, 3, 5):
        for i, t in enumerate(itertools.product("ATCG", repeat=k)):   # kmers.py:23 / conftest.py:201
            s = "".join(t)
            assert codec.encode(s, k) == i and codec.decode(i, k) == s
    assert codec.encode("ACGN", 4) is None and codec.encode("acgt", 4) is None and codec.encode("ACG", 4) is None
    s = "TCGCTTTCGTTCATG"
    assert codec.decode_many(codec.windows_to_codes(s, 11), 11) == [s[i:i + 11] for i in range(5)]


def test_edge_node_classes():
    from prairiedog_b200 import Edge, Node, concat_values
    e = Edge("A", "B", labels={"incr": "7", "x": 1})
    assert e.edge_value == 7 and e.labels == {"x": 1} and e.origin == "e"
    assert Edge("A", "B", edge_value="3").edge_value == 3
    assert concat_values((Node("ABC"), Node("BCD"), Node("CDE"))) == "ABCDE"
    assert concat_values(()) == ""


def test_matching_edges_rule():
    from prairiedog_b200 import Edge, Graph
    s = (Edge("a", "b", "T", 3), Edge("a", "c", "T", 9), Edge("a", "d", "U", 0))
    t = (Edge("x", "z", "T", 5),)
    ok, st = Graph.matching_edges(s, t)
    assert ok and [e.edge_value for e in st] == [3]
    assert Graph.matching_edges(s[1:], t) == (False, ())


def test_layout_follows_reference_loader():
    """fasta.Layout rebuilds Kmers._load's sequences from the device header table; here the table
    is derived from the golden header / sequence lists of the corner cases."""
    from prairiedog_b200 import fasta
    for name, rec in golden("corner_cases.json").items():
        if "raises" in rec or name == "lower_and_n":
            continue
        # residue index of each header = residues of all sequences flushed before it
        text = open(data_path(os.path.join("corner", name + ".fasta")), newline="").read()
        table, res = [], 0
        for ln in text.split("\n"):
            ln = ln.rstrip()
            if ln.startswith(">"):
                table.append((0, 0, res))
            else:
                res += len(ln)
        lay = fasta.Layout(np.array(table, dtype=np.uint64), res)
        assert lay.seq_len.tolist() == [len(s) for s in rec["sequences"]], name
        assert lay.aligned == (len(rec["headers"]) == len(rec["sequences"])), name
    lay = fasta.Layout(np.array([(0, 2, 0), (9, 11, 6), (14, 16, 8), (21, 23, 10)], dtype=np.uint64), 16)
    assert not lay.benign(4)          # two adjacent short contigs
    lay = fasta.Layout(np.array([(0, 2, 0), (9, 11, 6), (14, 16, 8)], dtype=np.uint64), 15)
    assert lay.benign(4) and lay.seq_len.tolist() == [6, 2, 7]


def test_integration_doc_names_every_entry_point():
    """INTEGRATION.md maps every exported symbol to the reference interface it replaces"""
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    missing = [n for n in header_symbols() if "`%s`" % n not in doc]
    assert missing == []
