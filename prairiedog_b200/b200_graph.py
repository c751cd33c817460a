"""B200Graph: the CUDA backend behind prairiedog's Graph plugin API.

Drop-in for LGGraph (prairiedog/lemon_graph.py:23-334): same methods, same return shapes, same
"not found is empty, not an exception" convention, data becomes visible to connected()/path() at
save() like a LemonGraph commit.  All graph content lives on the GPU (libpdgpu.so); this class only
interns strings (edge_type labels, node values that are not clean k-mers) and converts results back
into Node / Edge objects.  Bulk entry points (add_genome / add_fasta / path_many) move whole
genomes and query batches in one call; the per-edge methods keep the reference's call-by-call API.
"""
import bisect
import logging
import os
import sys
import typing

import numpy as np

from . import codec, fasta
from .edge import Edge
from .engine import Engine, HIT_EXPLICIT
from .errors import GraphException
from .graph import Graph
from .node import Node, DEFAULT_NODE_TYPE

log = logging.getLogger("prairiedog")

ET_DELIMITER = ' in '          # prairiedog/subgraph_ref.py:17
_FLUSH_AT = 1 << 16


class _File(object):
    """One bulk-loaded FASTA file: its contigs own the type ids [type_base, type_base + n)."""

    def __init__(self, basename, host, table, type_base, residue_base, seq_begin, seq_len):
        self.basename, self.host, self.table = basename, host, table
        self.type_base, self.n = type_base, int(table.shape[0])
        self.residue_base, self.seq_begin, self.seq_len = residue_base, seq_begin, seq_len
        self._strings = None

    def strings(self):
        if self._strings is None:
            self._strings = [h + ET_DELIMITER + self.basename for h in fasta.header_strings(self.host, self.table)]
        return self._strings


class B200Graph(Graph):
    def __init__(self, k: int = 11, device: int = 0, path_limit: int = None):
        self.k = k
        self.engine = Engine(k, device)
        # lemon_graph.py:229-236 skips a pair when its length exceeds sys.getrecursionlimit()
        self.path_limit = sys.getrecursionlimit() if path_limit is None else path_limit
        self._dist = None           # (rank, world, device) once distributed.attach() opened the communicator
        self._reset_host()

    def _reset_host(self):
        self._types = {}            # type id -> edge_type string (explicit / materialised)
        self._type_ids = {}         # edge_type string -> type id
        self._next_type = 0
        self._files = []            # bulk files, ascending type_base
        self._file_bases = []
        self._by_basename = {}
        self._explicit_suffixes = set()   # every text after an ' in ' of an explicit edge_type
        self._opaque = []           # opaque node values (not clean k-mers)
        self._opaque_ids = {}
        self._node_meta = {}        # node key -> (node_type, labels)
        self._edge_labels = {}      # (src key, tgt key, type id, value) -> labels
        self._graph_labels = {}
        self._st_src, self._st_tgt, self._st_type, self._st_val, self._st_nodes = [], [], [], [], []
        self._unsaved = False
        self._published = False     # nothing saved yet: queries see an empty graph
        self._type_bases = None     # multi-rank: global type numbering, agreed at the first query after a save()

    # ------------------------------------------------------------------ interning
    def _key(self, value, create=True):
        """Device key of a node value: the k-mer code, or an opaque id with bit 63 set.  Like LGGraph,
        whose store keeps str(value), a non-string value (the int labels of encode=True,
        subgraph_ref.py:54,65) reads back as its string."""
        if not isinstance(value, str):
            value = str(value)
        code = codec.encode(value, self.k)
        if code is not None:
            return code
        i = self._opaque_ids.get(value)
        if i is None:
            if not create:
                return None
            i = len(self._opaque)
            self._opaque.append(value)
            self._opaque_ids[value] = i
        return codec.OPAQUE_BIT | i

    def _value(self, key):
        key = int(key)
        if key & codec.OPAQUE_BIT:
            return self._opaque[key & ~codec.OPAQUE_BIT]
        return codec.decode(key, self.k)

    def _materialise(self, f):
        for i, s in enumerate(f.strings()):
            self._types[f.type_base + i] = s
            self._type_ids.setdefault(s, f.type_base + i)

    def _type_id(self, s):
        tid = self._type_ids.get(s)
        if tid is not None:
            return tid
        # could be the label of a bulk-loaded contig whose strings were never built
        pos = s.find(ET_DELIMITER)
        while pos >= 0:
            for f in self._by_basename.get(s[pos + len(ET_DELIMITER):], ()):
                self._materialise(f)
            pos = s.find(ET_DELIMITER, pos + 1)
        tid = self._type_ids.get(s)
        if tid is None:
            tid = self._next_type
            self._next_type += 1
            self._types[tid] = s
            self._type_ids[s] = tid
            pos = s.find(ET_DELIMITER)
            while pos >= 0:      # a later bulk load of a file with this name must compare labels
                self._explicit_suffixes.add(s[pos + len(ET_DELIMITER):])
                pos = s.find(ET_DELIMITER, pos + 1)
        return tid

    def _type_string(self, tid):
        tid = int(tid)
        s = self._types.get(tid)
        if s is None:
            f = self._files[bisect.bisect_right(self._file_bases, tid) - 1]
            s = f.strings()[tid - f.type_base]
        return s

    # ------------------------------------------------------------------ per-edge API
    def upsert_node(self, node: Node, echo: bool = True) -> typing.Optional[Node]:
        """lemon_graph.py:99-113."""
        key = self._key(node.value)
        self._st_nodes.append(key)
        meta = self._node_meta.get(key)
        if meta is None or node.labels is not None or node.node_type != DEFAULT_NODE_TYPE:
            labels = dict(meta[1]) if meta and meta[1] else {}
            if node.labels:
                labels.update(node.labels)
            self._node_meta[key] = (node.node_type, labels or None)
        self._unsaved = True
        if len(self._st_nodes) >= _FLUSH_AT:
            self._flush()
        if echo:
            return self._make_node(key)

    def add_edge(self, edge: Edge, echo: bool = True) -> typing.Optional[Edge]:
        """lemon_graph.py:115-128: both endpoints are upserted, the edge is unique on
        (type, value, src, tgt)."""
        ks, kt = self._key(edge.src), self._key(edge.tgt)
        tid = self._type_id(edge.edge_type)
        val = int(edge.edge_value)
        if not -(1 << 31) <= val < (1 << 31):
            raise ValueError("edge_value {} does not fit the 32-bit value column".format(val))
        self._st_src.append(ks)
        self._st_tgt.append(kt)
        self._st_type.append(tid)
        self._st_val.append(val)
        if edge.labels:
            self._edge_labels.setdefault((ks, kt, tid, val), {}).update(edge.labels)
        self._unsaved = True
        if len(self._st_src) >= _FLUSH_AT:
            self._flush()
        if echo:
            return Edge(src=ks + 1, tgt=kt + 1, edge_type=edge.edge_type, edge_value=val,
                        labels=self._edge_labels.get((ks, kt, tid, val)))

    def _flush(self):
        if self._st_src:
            self.engine.add_edges(np.array(self._st_src, dtype=np.uint64), np.array(self._st_tgt, dtype=np.uint64),
                                  np.array(self._st_type, dtype=np.uint32), np.array(self._st_val, dtype=np.int32))
            self._st_src, self._st_tgt, self._st_type, self._st_val = [], [], [], []
        if self._st_nodes:
            self.engine.add_nodes(np.array(self._st_nodes, dtype=np.uint64))
            self._st_nodes = []

    def save(self, f: str = None):
        """lemon_graph.py:154-157: commit.  Everything added so far becomes visible to queries.
        Like LGGraph.save the file name is ignored (SubgraphRef passes its .rdf name to every
        backend, subgraph_ref.py:117-125); snapshot(f) writes the graph to disk."""
        self._flush()
        self.engine.finalize()
        self._unsaved = False
        self._published = True
        self._type_bases = None

    def snapshot(self, f: str) -> str:
        """Publishes and writes the graph to ``f`` (the role of the LMDB file the Snakefile copies to
        outputs/pangenome_*.g, Snakefile:99-100); B200Graph.load reads it back."""
        from . import snapshot
        self._publish()
        return snapshot.save(self, f)

    @classmethod
    def load(cls, f: str, device: int = 0, path_limit: int = None) -> "B200Graph":
        """A graph from a file written by snapshot(f): the genome store and the contig table go back to
        the GPU and the index is rebuilt there."""
        from . import snapshot
        meta, _ = snapshot.read(f)
        g = cls(k=meta["k"], device=device, path_limit=path_limit)
        return snapshot.load(g, f)

    def _publish(self):
        if self._unsaved:
            self.save()

    def clear(self):
        self.engine.clear()
        self._reset_host()

    # ------------------------------------------------------------------ bulk API
    def add_fasta_bytes(self, name, host_bytes) -> int:
        """Loads one FASTA image (uint8 numpy array / bytes) whose edge labels use ``name`` as the
        file name.  Returns the number of k-mers it contributes (what update_graph returns,
        subgraph_ref.py:115).  Raises ValueError for inputs the bulk path cannot represent
        exactly; SubgraphRef.update_graph then uses the per-edge API instead."""
        host = np.frombuffer(host_bytes, dtype=np.uint8) if not isinstance(host_bytes, np.ndarray) else host_bytes
        d = self.engine.to_device(host)
        return self._add_device(name, host, d)

    def add_fasta(self, path) -> int:
        return self.add_fasta_bytes(os.path.basename(path), np.fromfile(path, dtype=np.uint8))

    def add_genome(self, km) -> int:
        """Bulk form of SubgraphRef.update_graph(km, ...) for a well-formed Kmers object."""
        if km.k != self.k:
            raise ValueError("Kmers k={} but graph k={}".format(km.k, self.k))
        if not km.well_formed:
            raise ValueError("{} needs the per-edge path".format(km))
        return self._add_device(str(km), km._host, km._d_bytes)

    def _add_device(self, name, host, d_bytes) -> int:
        info, table = self.engine.ingest(d_bytes)
        return self._register(name, host, int(info.n_dirty), int(info.appended), int(info.n_residues),
                              int(info.residue_base), table)

    def _register(self, name, host, n_dirty, appended, n_residues, residue_base, table) -> int:
        """Host bookkeeping of one parsed file: contig layout (kmers.py:61-75), type ids of its
        edge labels, contig declaration.  table: rows {byte_begin, byte_end, residue_index}."""
        if n_dirty != 0 or not appended:
            raise ValueError("{}: {} bytes outside the 2-bit alphabet".format(name, n_dirty))
        lay = fasta.Layout(table, n_residues)
        if not lay.benign(self.k):
            # residues stay in the store but no contig refers to them
            raise ValueError("{}: irregular FASTA layout needs the per-edge path".format(name))
        base = residue_base
        f = _File(name, host, table, self._next_type, base, lay.seq_begin, lay.seq_len)
        type_ids = np.arange(f.type_base, f.type_base + f.n, dtype=np.uint32)
        self._next_type += f.n
        if name in self._by_basename or name in self._explicit_suffixes or self._dup_headers(host, table):
            # the same label may occur twice: resolve strings now so that equal labels share an id
            # (libpdgpu then merges those contigs with set semantics)
            for other in self._by_basename.get(name, ()):
                self._materialise(other)
            for i, s in enumerate(f.strings()):
                tid = self._type_ids.setdefault(s, f.type_base + i)
                self._types.setdefault(tid, s)
                type_ids[i] = tid
        self._files.append(f)
        self._file_bases.append(f.type_base)
        self._by_basename.setdefault(name, []).append(f)
        self.engine.add_contigs(lay.seq_begin + base, lay.seq_len, type_ids)
        self._unsaved = True
        long_ = lay.seq_len > self.k
        return int((lay.seq_len[long_] - self.k + 1).sum())

    def add_fasta_batch(self, files) -> typing.List[int]:
        """Bulk form of the loop over genome files (Snakefile rule kmers + pangenome): ``files`` is
        a list of (name, host uint8 array, CUDA uint8 tensor or None).  With device tensors the
        images are parsed where they are; if any tensor is None the HOST arrays (pinned memory
        recommended) cross the C ABI and libpdgpu stages them itself, overlapping the copies with
        the parse.  Returns the k-mer count of every file (subgraph_ref.py:115).  Files the bulk
        path cannot represent exactly are skipped and reported through a ValueError whose
        ``failed`` attribute lists their indices (callers route those through the per-edge API)."""
        if not files:
            return []
        names = [f[0] for f in files]
        hosts = [f[1] if isinstance(f[1], np.ndarray) else np.frombuffer(f[1], dtype=np.uint8) for f in files]
        devs = [f[2] if len(f) > 2 else None for f in files]
        use_host = any(d is None for d in devs)
        if use_host:
            ptrs = [h.ctypes.data if h.size else 0 for h in hosts]
        else:
            ptrs = [d.data_ptr() if d.numel() else 0 for d in devs]
        infos, table, row_base = self.engine.ingest_batch(ptrs, [h.size for h in hosts], host=use_host)
        k = self.k
        n = len(files)
        nh = infos["n_headers"].astype(np.int64)
        nres = infos["n_residues"].astype(np.int64)
        rbase = infos["residue_base"].astype(np.int64)
        # fast path: every file clean, aligned (a header first, no empty record), every contig
        # longer than k, labels certainly distinct -> one vectorised pass for the whole batch
        fast = bool((infos["n_dirty"] == 0).all() and (infos["appended"] == 1).all() and (nh > 0).all()
                    and len(set(names)) == n
                    and not any(nm in self._by_basename or nm in self._explicit_suffixes for nm in names))
        if fast:
            res = table[:, 2].astype(np.int64)
            end = np.empty_like(res)
            end[:-1] = res[1:]
            end[row_base[1:] - 1] = nres
            lens = end - res
            file_of = np.repeat(np.arange(n, dtype=np.int64), nh)
            hashes = table[:, 3]
            mixed = hashes ^ (file_of.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15))
            fast = bool((res[row_base[:-1]] == 0).all() and (lens > k).all() and (hashes != 0).all()
                        and self._all_distinct(mixed))
        if fast:
            type0 = self._next_type
            counts = []
            kmers = np.add.reduceat(lens - (k - 1), row_base[:-1])
            for i in range(n):
                a, b = int(row_base[i]), int(row_base[i + 1])
                f = _File(names[i], hosts[i], table[a:b, :3], type0 + a, int(rbase[i]), res[a:b], lens[a:b])
                self._files.append(f)
                self._file_bases.append(f.type_base)
                self._by_basename.setdefault(names[i], []).append(f)
                counts.append(int(kmers[i]))
            self._next_type = type0 + int(row_base[-1])
            self.engine.add_contigs(res + rbase[file_of], lens,
                                    np.arange(type0, type0 + res.size, dtype=np.uint32))
            self._unsaved = True
            return counts
        counts, failed, msgs = [], [], []
        for i in range(n):
            a, b = int(row_base[i]), int(row_base[i + 1])
            try:
                counts.append(self._register(names[i], hosts[i], int(infos["n_dirty"][i]), int(infos["appended"][i]),
                                             int(nres[i]), int(rbase[i]), np.ascontiguousarray(table[a:b, :3])))
            except ValueError as e:
                counts.append(0)
                failed.append(i)
                msgs.append(str(e))
        if failed:
            err = ValueError("; ".join(msgs))
            err.failed = failed
            err.counts = counts
            raise err
        return counts

    @staticmethod
    def _all_distinct(keys):
        """np.unique hashes / sorts with an inverse pass; a plain sort is ~15x faster at 20k labels"""
        if keys.size < 2:
            return True
        srt = np.sort(keys)
        return bool((srt[1:] != srt[:-1]).all())

    @staticmethod
    def _dup_headers(host, table):
        """True when two header lines of the file give the same label.  Labels are the RIGHT-STRIPPED lines
        (kmers.py:63 `line.rstrip()`), so '>c1' and '>c1 ' are one edge_type; bytes.rstrip() strips the same ASCII
        whitespace set as the device hash of the batched path (csrc/ingest_batch.cu fb_hash_kernel)."""
        if table.shape[0] < 2:
            return False
        mv = memoryview(host)
        seen = set()
        for b, e in table[:, :2].tolist():
            seen.add(bytes(mv[b:e]).rstrip())
        return len(seen) != table.shape[0]

    # ------------------------------------------------------------------ enumeration
    def _make_node(self, key) -> Node:
        key = int(key)
        ntype, labels = self._node_meta.get(key, (DEFAULT_NODE_TYPE, None))
        return Node(value=self._value(key), node_type=ntype, db_id=key + 1, labels=labels)

    @property
    def nodes(self) -> typing.Set[Node]:
        """lemon_graph.py:133-140.  db_id = node key + 1 (for a k-mer: its rank among all 4^k
        k-mers in 'ATCG' product order, plus one -- the reference's kmer_map fixture)."""
        self._publish()
        return set(self._make_node(c) for c in self.engine.export_nodes().tolist())

    @property
    def edges(self) -> typing.Set[Edge]:
        """lemon_graph.py:142-149: src / tgt are the numeric db ids of the endpoint nodes."""
        self._publish()
        src, tgt, ty, val = self.engine.export_edges()
        out = set()
        for s, t, y, v in zip(src.tolist(), tgt.tolist(), ty.tolist(), val.tolist()):
            out.add(Edge(src=s + 1, tgt=t + 1, edge_type=self._type_string(y), edge_value=v,
                         labels=self._edge_labels.get((s, t, y, v))))
        return out

    def canonical(self):
        """(sorted node values, sorted (src value, tgt value, edge_type, edge_value)) -- the
        canonical forms of SURVEY.md A.2 used by the parity tests."""
        self._publish()
        nodes = sorted(self._value(c) for c in self.engine.export_nodes().tolist())
        src, tgt, ty, val = self.engine.export_edges()
        edges = sorted((self._value(s), self._value(t), self._type_string(y), v)
                       for s, t, y, v in zip(src.tolist(), tgt.tolist(), ty.tolist(), val.tolist()))
        return nodes, edges

    @property
    def edgelist(self) -> typing.Generator:
        self._publish()
        src, tgt, ty, val = self.engine.export_edges()
        for s, t, y, v in zip(src.tolist(), tgt.tolist(), ty.tolist(), val.tolist()):
            yield self._value(s), self._value(t), {'edge_type': self._type_string(y), 'edge_value': v}

    def get_labels(self, node: str) -> dict:
        key = self._key(node, create=False)
        ntype, labels = self._node_meta.get(key, (DEFAULT_NODE_TYPE, None))
        d = {'type': ntype, 'value': node}
        d.update(labels or {})
        return d

    def set_graph_labels(self, labels: dict):
        self._graph_labels.update(labels)

    def filter(self):
        """Graph.filter() (graph.py:62-63) as the one backend that implements it defines it
        (networkx_graph.py:53-75; LGGraph.filter is `pass`): on the undirected simple graph of the node pairs, drop
        the nodes of degree <= int(0.9 * max degree), then the isolates; the nodes go together with every labelled
        edge that touches them.  Degrees, both removals and the surviving edge list are computed on the device
        (pdg_filter_*); the graph is then rebuilt from the surviving labelled edges through the per-edge path, so
        afterwards it holds exactly the induced sub-multigraph (committed)."""
        if self._dist is not None:
            raise NotImplementedError("filter() of a multi-rank graph: a node's edges are spread over the ranks")
        self._publish()
        if not self._published:
            return
        n_nodes, src, tgt, ty, val = self.engine.filter_edges()
        self.engine.clear()            # the store goes too: what is left is explicit edges (labels stay on the host)
        self._st_src, self._st_tgt, self._st_type, self._st_val, self._st_nodes = [], [], [], [], []
        if src.size:
            self.engine.add_edges(src, tgt, ty, val)
        self.engine.finalize()
        self._unsaved = False
        self._published = True

    def __len__(self):
        self._publish()
        return self.engine.counts()[0]

    # ------------------------------------------------------------------ queries
    def _codes(self, values):
        """Keys of query node values; unknown opaque values map to a key no node has."""
        missing = codec.OPAQUE_BIT | ((1 << 62) - 1)
        return np.array([k if (k := self._key(v, create=False)) is not None else missing for v in values],
                        dtype=np.uint64)

    def connected_many(self, pairs):
        """Batched connected(): list of (bool, tuple[Edge]) in the order of ``pairs``.  Multi-rank: collective, the
        start edges of every rank's shard in rank order."""
        out = self._connected_many_local(pairs)
        if self._dist is not None:
            from . import distributed
            parts = distributed.gather_objects(out)
            out = []
            for q in range(len(pairs)):
                edges = sum((tuple(p[q][1]) for p in parts), ())
                out.append((len(edges) > 0, edges))
        return out

    def _connected_many_local(self, pairs):
        a = self._codes([p[0] for p in pairs])
        b = self._codes([p[1] for p in pairs])
        if not self._published:
            return [(False, ())] * len(pairs)
        off, hits = self.engine.query_connected(a, b)
        out = []
        for q in range(len(pairs)):
            hs = hits[int(off[q]):int(off[q + 1])]
            out.append((len(hs) > 0, tuple(self._start_edge(int(a[q]), h) for h in hs)))
        return out

    def _start_edge(self, src_key, h) -> Edge:
        tid, i = int(h["type_id"]), int(h["i"])
        codes, _ = self.engine.walk(src_key, {"type_id": tid, "i": i, "j": i, "aux": h["aux"]})
        tgt_key = int(codes[1])
        return Edge(src=src_key + 1, tgt=tgt_key + 1, edge_type=self._type_string(tid), edge_value=i,
                    labels=self._edge_labels.get((src_key, tgt_key, tid, i)))

    def connected(self, node_a: str, node_b: str) -> typing.Tuple[bool, typing.Tuple]:
        """lemon_graph.py:178-219."""
        return self.connected_many([(node_a, node_b)])[0]

    def path_many(self, pairs, limit=None, local=False):
        """Batched path(): (codes of the a's, offsets[n_pairs + 1], hit records {type_id, i, j, aux}) -- no Node
        objects are built.  In a multi-rank job (distributed.attach) the call is COLLECTIVE: every rank passes the
        same pairs, runs them against its own shard and receives the hits of the whole pangenome, type ids in the
        global numbering (``self.hit_owner`` then names the rank that holds each hit's contig; the returned
        arrays are views of a pinned buffer that the next collective call reuses); ``local=True`` keeps the answer of
        this rank's shard."""
        a = self._codes([p[0] for p in pairs])
        b = self._codes([p[1] for p in pairs])
        lim = self.path_limit if limit is None else limit
        self.hit_owner = None
        if self._dist is not None and not local:
            from . import distributed
            rank, world, dev = self._dist
            if self._type_bases is None:
                self._type_bases = distributed.type_bases(self._next_type, dev)
            base = int(self._type_bases[rank])
            if self._published and dev is not None and getattr(dev, "type", "") == "cuda":
                # the hits stay on the device from the query kernels to the merged answer
                off_t, hits_t = self.engine.query_paths_device(a, b, lim)
                off, hits, self.hit_owner = distributed.exchange_hits_device(off_t, hits_t, base)
                return a, off, hits
            from .engine import HIT_DTYPE
            if self._published:
                off, hits = self.engine.query_paths(a, b, lim)
            else:
                off, hits = np.zeros(len(pairs) + 1, dtype=np.uint64), np.zeros(0, dtype=HIT_DTYPE)
            off, hits, self.hit_owner = distributed.exchange_hits(off, hits, base, dev)
            return a, off, hits
        if not self._published:
            from .engine import HIT_DTYPE
            return a, np.zeros(len(pairs) + 1, dtype=np.uint64), np.zeros(0, dtype=HIT_DTYPE)
        off, hits = self.engine.query_paths(a, b, lim)
        return a, off, hits

    def query_many(self, pairs, limit=None, strings=True):
        """Prairiedog.query (prairiedog.py:17-33) for a batch of (src, dst) pairs: per pair the list of
        hit dicts {'string', 'edge_type', 'i', 'j'}; 'string' is concat_values (node.py:22-37) of the
        hit's node chain, spelled from the codes without building Node objects."""
        if self._dist is not None:
            return self._query_many_distributed(pairs, limit, strings)
        a, off, hits = self.path_many(pairs, limit)
        spell = self._hit_strings(a, off, hits) if strings else None
        out = []
        tid, hi, hj = hits["type_id"].tolist(), hits["i"].tolist(), hits["j"].tolist()
        for q in range(len(pairs)):
            row = []
            for x in range(int(off[q]), int(off[q + 1])):
                d = {'edge_type': self._type_string(tid[x]), 'i': hi[x], 'j': hj[x]}
                if strings:
                    d['string'] = spell[x]
                row.append(d)
            out.append(row)
        return out

    def _query_many_distributed(self, pairs, limit, strings):
        """query_many over an N-rank graph (collective): every rank spells the hits of ITS shard -- the labels and
        the residues live there -- and the rows are concatenated in rank order, which is the order a single-process
        graph returns them in."""
        from . import distributed
        a, off, hits = self.path_many(pairs, limit, local=True)
        spell = self._hit_strings(a, off, hits) if strings else None
        rank, world, dev = self._dist
        if self._type_bases is None:
            self._type_bases = distributed.type_bases(self._next_type, dev)
        tid, hi, hj = hits["type_id"].tolist(), hits["i"].tolist(), hits["j"].tolist()
        mine = []
        for q in range(len(pairs)):
            row = []
            for x in range(int(off[q]), int(off[q + 1])):
                d = {'edge_type': self._type_string(tid[x]), 'i': hi[x], 'j': hj[x]}
                if strings:
                    d['string'] = spell[x]
                row.append(d)
            mine.append(row)
        parts = distributed.gather_objects(mine)
        return [sum((parts[r][q] for r in range(world)), []) for q in range(len(pairs))]

    def _hit_strings(self, a, off, hits):
        """concat_values (node.py:22-37) of every hit of the LAST path_many call: genome-part hits are decoded on
        the device in one pdg_query_strings call (contig[i : j + 1 + k]); explicit-part hits, whose node values can
        be opaque host strings, follow their chain with pdg_query_walk (lemon_graph.py:244-264)."""
        n = int(hits.size)
        if n == 0:
            return []
        soff, sbytes = self.engine.query_strings()
        raw = sbytes.tobytes()
        lens = (hits["j"].astype(np.int64) - hits["i"].astype(np.int64) + 1 + self.k).tolist()
        explicit = ((hits["aux"] & HIT_EXPLICIT) != 0).tolist()
        soff = soff.tolist()
        out = [None] * n
        qof = np.repeat(np.arange(len(off) - 1), np.diff(off.astype(np.int64))).tolist() if any(explicit) else None
        for x in range(n):
            if explicit[x]:
                codes, status = self.engine.walk(int(a[qof[x]]), hits[x])
                # lemon_graph.py:254-258: exactly one chain must spell the path (an AssertionError there too)
                assert status == 0, "Length of chains is {} != 1".format(0 if status == 1 else 2)
                vals = [self._value(c) for c in codes.tolist()]
                out[x] = vals[0] + ''.join(v[-1] for v in vals[1:])
            else:
                out[x] = raw[soff[x]:soff[x] + lens[x]].decode("ascii")
        return out

    def path(self, node_a: str, node_b: str) -> typing.Tuple[tuple, tuple]:
        """lemon_graph.py:275-334.  Multi-rank: collective, the paths of every rank's shard in rank order."""
        paths, metas = self._path_local(node_a, node_b)
        if self._dist is not None:
            from . import distributed
            parts = distributed.gather_objects((paths, metas))
            paths = sum((tuple(p[0]) for p in parts), ())
            metas = sum((tuple(p[1]) for p in parts), ())
        return paths, metas

    def _path_local(self, node_a, node_b):
        a, off, hits = self.path_many([(node_a, node_b)], local=True)
        paths, metas = [], []
        spell = self._hit_strings(a, off, hits)
        k = self.k
        for x, h in enumerate(hits):
            if int(h["aux"]) & HIT_EXPLICIT:
                codes, status = self.engine.walk(int(a[0]), h)
                # lemon_graph.py:254-258: exactly one chain must spell the path (an AssertionError there too)
                assert status == 0, "Length of chains is {} != 1".format(0 if status == 1 else 2)
                codes = codes.tolist()
            else:       # genome part: node t of the path is the k-mer at offset t of the hit's string
                s = spell[x]
                codes = [codec.encode(s[t:t + k], k) for t in range(len(s) - k + 1)]
            if len(codes) < 2:          # lemon_graph.py:315-316
                raise GraphException(self)
            paths.append(tuple(self._make_node(c) for c in codes))
            meta = {'edge_type': self._type_string(h["type_id"])}
            lab = self._edge_labels.get((int(codes[0]), int(codes[1]), int(h["type_id"]), int(h["i"])))
            if lab:
                meta.update(lab)
            metas.append(meta)
        return tuple(paths), tuple(metas)
