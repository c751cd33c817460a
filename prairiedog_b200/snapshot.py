"""Persistent snapshot of a B200Graph and exporters to the reference's interchange formats
(SURVEY.md 8f rank 3).

The reference persists a pangenome as the LMDB file behind LGGraph (lemon_graph.py:154-157; the
Snakefile copies it to outputs/pangenome_*.g, Snakefile:99-100), as a networkx gpickle
(networkx_graph.py:41-43) or as Dgraph RDF N-Quads (dgraph.py:576-613).  Here the snapshot holds the
INPUT of pdg_finalize -- the 2-bit genome store, the declared contigs, the explicit edges / nodes --
plus the host string tables; load() puts them back on the GPU and rebuilds the index there, which
is faster than reading a 4 B/edge index from disk.

File layout (little endian): magic b"PDGSNAP1", u64 header length, JSON header, then the arrays,
each 64-byte aligned, at the offsets the header lists.
"""
import hashlib
import json
import os

import numpy as np

from .node import DEFAULT_NODE_TYPE

MAGIC = b"PDGSNAP1"
DEFAULT_EDGE_PREDICATE = "fd"      # prairiedog/dgraph.py:23
_ARRAYS = ("words", "c_begin", "c_len", "c_type", "x_src", "x_tgt", "x_type", "x_val", "x_nodes")


def _align(n, a=64):
    return (n + a - 1) // a * a


def save(graph, path):
    """Writes the published state of ``graph`` (a B200Graph) to ``path``."""
    graph._flush()
    snap = graph.engine.snapshot_export()
    files = []
    for f in graph._files:
        files.append({"basename": f.basename, "type_base": int(f.type_base), "n": int(f.n),
                      "residue_base": int(f.residue_base), "strings": f.strings(),
                      "seq_begin": [int(x) for x in f.seq_begin], "seq_len": [int(x) for x in f.seq_len]})
    meta = {
        "k": graph.k, "n_residues": snap["n_residues"], "next_type": graph._next_type,
        "types": [[int(t), s] for t, s in graph._types.items()],
        "files": files, "opaque": list(graph._opaque),
        "explicit_suffixes": sorted(graph._explicit_suffixes),
        "node_meta": [[int(key), nt, labels] for key, (nt, labels) in graph._node_meta.items()],
        "edge_labels": [[int(a), int(b), int(t), int(v), labels] for (a, b, t, v), labels in graph._edge_labels.items()],
        "graph_labels": graph._graph_labels,
        "arrays": {},
    }
    off = 0
    for name in _ARRAYS:
        a = np.ascontiguousarray(snap[name])
        meta["arrays"][name] = {"dtype": a.dtype.str, "count": int(a.size), "offset": off}
        off = _align(off + a.nbytes)
    header = json.dumps(meta).encode("utf-8")
    data0 = _align(len(MAGIC) + 8 + len(header))
    parent = os.path.dirname(os.path.abspath(path))
    os.makedirs(parent, exist_ok=True)
    tmp = path + ".tmp"
    with open(tmp, "wb") as out:
        out.write(MAGIC)
        out.write(np.uint64(len(header)).tobytes())
        out.write(header)
        out.write(b"\0" * (data0 - out.tell()))
        for name in _ARRAYS:
            a = np.ascontiguousarray(snap[name])
            out.seek(data0 + meta["arrays"][name]["offset"])
            out.write(a.tobytes())
    os.replace(tmp, path)     # a reader never sees a half-written snapshot
    return path


def read(path):
    """(meta, arrays) of a snapshot file; arrays are numpy views of one read of the file."""
    with open(path, "rb") as f:
        blob = f.read()
    if blob[:8] != MAGIC:
        raise ValueError("{} is not a prairiedog_b200 snapshot".format(path))
    hlen = int(np.frombuffer(blob, dtype=np.uint64, count=1, offset=8)[0])
    meta = json.loads(blob[16:16 + hlen].decode("utf-8"))
    data0 = _align(16 + hlen)
    arrays = {}
    for name, d in meta["arrays"].items():
        if d["count"] == 0:       # may sit at (or past) the end of the file
            arrays[name] = np.zeros(0, dtype=np.dtype(d["dtype"]))
            continue
        arrays[name] = np.frombuffer(blob, dtype=np.dtype(d["dtype"]), count=d["count"], offset=data0 + d["offset"])
    return meta, arrays


def load(graph, path):
    """Replaces the content of ``graph`` by the snapshot and publishes it (index rebuilt on the GPU)."""
    from .b200_graph import _File
    meta, arr = read(path)
    if meta["k"] != graph.k:
        raise ValueError("snapshot has k={}, graph has k={}".format(meta["k"], graph.k))
    graph.clear()
    graph.engine.store_import(arr["words"], meta["n_residues"])
    if arr["c_begin"].size:
        graph.engine.add_contigs(arr["c_begin"], arr["c_len"], arr["c_type"])
    if arr["x_src"].size:
        graph.engine.add_edges(arr["x_src"], arr["x_tgt"], arr["x_type"], arr["x_val"])
    if arr["x_nodes"].size:
        graph.engine.add_nodes(arr["x_nodes"])
    graph._next_type = meta["next_type"]
    for t, s in meta["types"]:
        graph._types[t] = s
        graph._type_ids.setdefault(s, t)
    for d in meta["files"]:
        f = _File(d["basename"], None, np.zeros((d["n"], 3), dtype=np.uint64), d["type_base"], d["residue_base"],
                  np.array(d["seq_begin"], dtype=np.int64), np.array(d["seq_len"], dtype=np.int64))
        f._strings = list(d["strings"])
        graph._files.append(f)
        graph._file_bases.append(f.type_base)
        graph._by_basename.setdefault(f.basename, []).append(f)
    graph._opaque = list(meta["opaque"])
    graph._opaque_ids = {v: i for i, v in enumerate(graph._opaque)}
    graph._explicit_suffixes = set(meta["explicit_suffixes"])
    graph._node_meta = {int(key): (nt, labels) for key, nt, labels in meta["node_meta"]}
    graph._edge_labels = {(a, b, t, v): labels for a, b, t, v, labels in meta["edge_labels"]}
    graph._graph_labels = dict(meta["graph_labels"])
    graph._unsaved = True
    graph.save()
    return graph


# ---------------------------------------------------------------------------- exporters
def write_edgelist(graph, path, data=False):
    """networkx edgelist lines 'src tgt' (networkx_graph.py:45-47 uses generate_edgelist(data=False));
    with data=True the attribute dict of the reference's add_edge (networkx_graph.py:22-25) follows."""
    n = 0
    with open(path, "w") as out:
        for s, t, d in graph.edgelist:
            out.write("{} {} {}\n".format(s, t, d) if data else "{} {}\n".format(s, t))
            n += 1
    return n


def rdf_lines(src, tgt, edge_type, edge_value):
    """The four N-Quads DgraphBulk.add_edge emits per edge (dgraph.py:576-589): an intermediate blank
    node named by sha1(src tgt type value) carries the label."""
    e = hashlib.sha1("{}{}{}{}".format(src, tgt, edge_type, edge_value).encode("utf-8")).hexdigest()
    p = DEFAULT_EDGE_PREDICATE
    return ('_:{} <{}> _:{} .'.format(src, p, e), '_:{} <{}> _:{} .'.format(e, p, tgt),
            '_:{} <type> "{}" .'.format(e, edge_type), '_:{} <value> "{}" .'.format(e, edge_value))


def write_rdf(graph, path, preload_nodes=True):
    """Dgraph bulk-loader input: one '_:{kmer} <km> "{kmer}" .' line per node (DgraphBulk.preload,
    dgraph.py:568-574, restricted to the k-mers present) and the four edge lines of rdf_lines."""
    n = 0
    with open(path, "w") as out:
        if preload_nodes:
            graph._publish()
            for c in graph.engine.export_nodes().tolist():
                v = graph._value(c)
                out.write('_:{v} <{nt}> "{v}" .\n'.format(v=v, nt=DEFAULT_NODE_TYPE))
        for s, t, d in graph.edgelist:
            for ln in rdf_lines(s, t, d['edge_type'], d['edge_value']):
                out.write(ln + "\n")
            n += 1
    return n
