"""Engine: one pdg_graph handle of libpdgpu.so with numpy-friendly wrappers.

Host arrays are numpy, device buffers are torch tensors whose ``data_ptr()`` crosses the C ABI
(include/pdgpu.h).  There is no CPU implementation: constructing an Engine without a CUDA device
raises PdgError.
"""
import ctypes

import numpy as np

from . import _ffi
from ._ffi import check

HIT_DTYPE = np.dtype([("type_id", np.uint32), ("i", np.int32), ("j", np.int32), ("aux", np.uint32)])
HIT_EXPLICIT = 0x80000000


def _u64(a):
    return np.ascontiguousarray(a, dtype=np.uint64)


class Engine(object):
    def __init__(self, k=11, device=0):
        self.lib = _ffi.load()
        self.k = int(k)
        self.device = int(device)
        self.h = self.lib.pdg_create(self.device, self.k)
        if not self.h:
            raise _ffi.PdgError(_ffi.PDG_E_CUDA, _ffi.last_error())

    def close(self):
        if getattr(self, "h", None):
            self.lib.pdg_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- ingest ----
    def to_device(self, host_bytes):
        """uint8 numpy array / bytes -> torch uint8 CUDA tensor (16-byte aligned by the allocator)."""
        import torch
        arr = np.frombuffer(host_bytes, dtype=np.uint8) if not isinstance(host_bytes, np.ndarray) else host_bytes
        if arr.size == 0:
            return torch.empty(0, dtype=torch.uint8, device="cuda:%d" % self.device)
        t = torch.from_numpy(np.ascontiguousarray(arr))
        return t.to("cuda:%d" % self.device)

    def ingest(self, d_bytes, n_bytes=None):
        """pdg_ingest_fasta on a CUDA uint8 tensor.  Returns (info, header_table[n_headers,3])."""
        n = int(d_bytes.numel()) if n_bytes is None else int(n_bytes)
        info = _ffi.IngestInfo()
        cap = 1024
        while True:
            table = np.zeros((cap, 3), dtype=np.uint64)
            rc = self.lib.pdg_ingest_fasta(self.h, d_bytes.data_ptr() if n else None, n,
                                           table.ctypes.data, cap, ctypes.byref(info))
            if rc == _ffi.PDG_E_CAPACITY:
                cap = int(info.n_headers)
                continue
            check(rc)
            return info, table[:int(info.n_headers)]

    def ingest_batch(self, ptrs, sizes, host=False):
        """pdg_ingest_fasta_batch.  ptrs: raw pointers (device, or host when ``host``) of the file
        images, sizes: their byte counts.  Returns (infos, table[n_rows, 4], row_base[n_files+1]):
        infos is a numpy structured array with the fields of pdg_ingest_info, file f owns rows
        table[row_base[f]:row_base[f+1]] = {byte_begin, byte_end, residue_index, hash}."""
        n = len(ptrs)
        p = (ctypes.c_void_p * max(n, 1))(*[int(x) if x else None for x in ptrs])
        sz = np.ascontiguousarray(sizes, dtype=np.uint64)
        infos = np.zeros(n, dtype=_ffi.INGEST_INFO_DTYPE)
        cap = max(256 * n, 1024) if getattr(self, "_batch_rows", 0) == 0 else self._batch_rows
        while True:
            table = np.empty((cap, 4), dtype=np.uint64)
            rc = self.lib.pdg_ingest_fasta_batch(self.h, p, sz.ctypes.data, n, _ffi.PDG_IN_HOST if host else 0,
                                                 table.ctypes.data, cap, infos.ctypes.data)
            if rc == _ffi.PDG_E_CAPACITY:
                cap = int(infos["n_headers"].sum())
                continue
            check(rc)
            break
        row_base = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(infos["n_headers"].astype(np.int64), out=row_base[1:])
        self._batch_rows = max(int(row_base[-1]), 1)
        return infos, table[:int(row_base[-1])], row_base

    def add_contigs(self, begin, length, type_id):
        begin, length = _u64(begin), _u64(length)
        type_id = np.ascontiguousarray(type_id, dtype=np.uint32)
        check(self.lib.pdg_add_contigs(self.h, begin.ctypes.data, length.ctypes.data, type_id.ctypes.data,
                                       begin.size))

    def kmer_stats(self, begin, length):
        begin, length = _u64(begin), _u64(length)
        n, u = ctypes.c_uint64(), ctypes.c_uint64()
        check(self.lib.pdg_kmer_stats(self.h, begin.ctypes.data, length.ctypes.data, begin.size,
                                      ctypes.byref(n), ctypes.byref(u)))
        return n.value, u.value

    def fetch_residues(self, begin, length):
        out = np.zeros(int(length), dtype=np.uint8)
        check(self.lib.pdg_fetch_residues(self.h, int(begin), int(length), out.ctypes.data))
        return out.tobytes().decode("ascii")

    def emit_codes(self, begin, length):
        import torch
        n = max(int(length) - self.k + 1, 0)
        d = torch.zeros(max(n, 1), dtype=torch.int64, device="cuda:%d" % self.device)
        check(self.lib.pdg_emit_codes(self.h, int(begin), int(length), d.data_ptr()))
        return d[:n].cpu().numpy().view(np.uint64)

    # ---- snapshot (pdg_snapshot_*, pdg_store_import) ----
    def snapshot_export(self):
        """dict of host arrays: the input of pdg_finalize (2-bit store, contigs, explicit part)"""
        sz = (ctypes.c_uint64 * 4)()
        check(self.lib.pdg_snapshot_sizes(self.h, sz))
        n_res, n_c, n_e, n_n = (int(x) for x in sz)
        words = np.zeros(max((n_res + 31) // 32, 1), dtype=np.uint64)
        out = {"words": words, "c_begin": np.zeros(n_c, np.uint64), "c_len": np.zeros(n_c, np.uint64),
               "c_type": np.zeros(n_c, np.uint32), "x_src": np.zeros(n_e, np.uint64), "x_tgt": np.zeros(n_e, np.uint64),
               "x_type": np.zeros(n_e, np.uint32), "x_val": np.zeros(n_e, np.int32), "x_nodes": np.zeros(n_n, np.uint64)}
        check(self.lib.pdg_snapshot_export(self.h, words.ctypes.data, words.size, *(
            out[f].ctypes.data for f in ("c_begin", "c_len", "c_type", "x_src", "x_tgt", "x_type", "x_val", "x_nodes"))))
        out["words"] = words[:(n_res + 31) // 32]
        out["n_residues"] = n_res
        return out

    def store_import(self, words, n_residues):
        words = _u64(words)
        if words.size < (int(n_residues) + 31) // 32:
            raise ValueError("store of %d residues needs %d words, got %d" % (n_residues, (n_residues + 31) // 32, words.size))
        check(self.lib.pdg_store_import(self.h, words.ctypes.data, int(n_residues)))

    # ---- per-edge API ----
    def add_edges(self, src, tgt, type_id, value):
        src, tgt = _u64(src), _u64(tgt)
        type_id = np.ascontiguousarray(type_id, dtype=np.uint32)
        value = np.ascontiguousarray(value, dtype=np.int32)
        check(self.lib.pdg_add_edges_host(self.h, src.ctypes.data, tgt.ctypes.data, type_id.ctypes.data,
                                          value.ctypes.data, src.size))

    def add_nodes(self, codes):
        codes = _u64(codes)
        check(self.lib.pdg_add_nodes_host(self.h, codes.ctypes.data, codes.size))

    # ---- build ----
    def finalize(self, flags=0):
        check(self.lib.pdg_finalize(self.h, int(flags)))

    def clear(self):
        check(self.lib.pdg_clear(self.h))

    def counts(self):
        a, b, c = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64()
        check(self.lib.pdg_counts(self.h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)))
        return a.value, b.value, c.value

    def stats(self):
        st = _ffi.BuildStats()
        check(self.lib.pdg_get_build_stats(self.h, ctypes.byref(st)))
        out = {f: getattr(st, f) for f, _ in st._fields_ if not f.startswith("pad")}
        out["ms_pass"] = list(out["ms_pass"])
        return out

    def export_nodes(self):
        n = self.counts()[0]
        out = np.zeros(n, dtype=np.uint64)
        check(self.lib.pdg_export_nodes(self.h, out.ctypes.data))
        return out

    def export_edges(self):
        n = self.counts()[1]
        src, tgt = np.zeros(n, dtype=np.uint64), np.zeros(n, dtype=np.uint64)
        ty, val = np.zeros(n, dtype=np.uint32), np.zeros(n, dtype=np.int32)
        check(self.lib.pdg_export_edges(self.h, src.ctypes.data, tgt.ctypes.data, ty.ctypes.data, val.ctypes.data))
        return src, tgt, ty, val

    def node_ids(self, codes):
        codes = _u64(codes)
        out = np.zeros(codes.size, dtype=np.uint64)
        check(self.lib.pdg_node_ids(self.h, codes.ctypes.data, codes.size, out.ctypes.data))
        return out

    def export_index(self):
        r, p = ctypes.c_uint64(), ctypes.c_uint64()
        check(self.lib.pdg_index_size(self.h, ctypes.byref(r), ctypes.byref(p)))
        codes = np.zeros(r.value, dtype=np.uint64)
        rowptr = np.zeros(r.value + 1, dtype=np.uint32)
        pos = np.zeros(p.value, dtype=np.uint32)
        check(self.lib.pdg_export_index(self.h, codes.ctypes.data, rowptr.ctypes.data, pos.ctypes.data))
        return codes, rowptr, pos

    def checksum(self):
        a, b = ctypes.c_uint64(), ctypes.c_uint64()
        check(self.lib.pdg_checksum(self.h, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    # ---- queries ----
    def _query(self, fn, a, b, *extra):
        a, b = _u64(a), _u64(b)
        tot = ctypes.c_uint64()
        check(fn(self.h, a.ctypes.data, b.ctypes.data, a.size, *extra, ctypes.byref(tot)))
        self._last_hits = tot.value
        off = np.zeros(a.size + 1, dtype=np.uint64)
        hits = np.zeros(tot.value, dtype=HIT_DTYPE)
        check(self.lib.pdg_query_fetch(self.h, off.ctypes.data, hits.ctypes.data if tot.value else None))
        return off, hits

    def query_paths_device(self, a, b, max_len=-1):
        """query_paths with the result left on the device: (offsets int64[n_q + 1], hits int32[n_hits, 4]) as torch
        CUDA tensors (pdg_query_fetch into device memory) -- the multi-rank exchange starts from these."""
        import torch
        a, b = _u64(a), _u64(b)
        tot = ctypes.c_uint64()
        check(self.lib.pdg_query_paths(self.h, a.ctypes.data, b.ctypes.data, a.size, int(max_len), ctypes.byref(tot)))
        self._last_hits = tot.value
        dev = "cuda:%d" % self.device
        off = torch.empty(a.size + 1, dtype=torch.int64, device=dev)
        hits = torch.empty((max(tot.value, 1), 4), dtype=torch.int32, device=dev)
        check(self.lib.pdg_query_fetch(self.h, off.data_ptr(), hits.data_ptr() if tot.value else None))
        return off, hits[:tot.value]

    def query_connected(self, a, b):
        """(offsets[n_q+1], hits): one hit per start edge (type_id, i)."""
        return self._query(self.lib.pdg_query_connected, a, b)

    def query_paths(self, a, b, max_len=-1):
        """(offsets[n_q+1], hits): one hit per (type_id, i, j) with j - i <= max_len."""
        return self._query(self.lib.pdg_query_paths, a, b, int(max_len))

    def walk(self, src_code, hit):
        """Node codes of one hit's path and the chain status (0 ok, 1 broken, 2 ambiguous)."""
        h = _ffi.Hit(int(hit["type_id"]), int(hit["i"]), int(hit["j"]), int(hit["aux"]))
        n = int(hit["j"]) - int(hit["i"]) + 2
        out = np.zeros(n, dtype=np.uint64)
        st = ctypes.c_int32()
        check(self.lib.pdg_query_walk(self.h, int(src_code), ctypes.byref(h), out.ctypes.data, ctypes.byref(st)))
        return out, st.value

    def query_strings(self):
        """Strings of every hit of the last query_paths call (pdg_query_strings_*): (offsets[n_hits+1] in bytes,
        uint8 array).  Hit h's string is bytes[off[h] : off[h] + (j - i + 1 + k)]; explicit-part hits have none."""
        import torch
        tot = ctypes.c_uint64()
        check(self.lib.pdg_query_strings_size(self.h, ctypes.byref(tot)))
        n = int(self._last_hits)
        off = np.zeros(n + 1, dtype=np.uint64)
        # pinned: one cudaMemcpyAsync straight into it.  The buffer is kept and only grows: cudaHostAlloc of a fresh
        # 1.5 GB buffer per call cost 0.7 s (2 GB/s end to end) against 30 ms for the copy itself.  The returned array
        # is a VIEW of it, valid until the next query_strings call.
        buf = getattr(self, "_pinned_strings", None)
        if buf is None or buf.numel() < tot.value:
            self._pinned_strings = buf = torch.empty(max(int(tot.value * 1.25), 1 << 20), dtype=torch.uint8,
                                                     pin_memory=True)
        check(self.lib.pdg_query_strings_fetch(self.h, off.ctypes.data, buf.data_ptr() if tot.value else None))
        return off, buf.numpy()[:tot.value]

    # ---- Graph.filter() ----
    def filter_edges(self, bottom_percentile=0.9):
        """networkx_graph.py:53-75 on the device: (n_kept_nodes, src, tgt, type, value) of the labelled edges between
        the nodes that survive the degree filter and the isolate filter.  The threshold is the reference's own
        expression, int(bottom_percentile * max_degrees), evaluated here in Python."""
        md = ctypes.c_uint32()
        check(self.lib.pdg_filter_degrees(self.h, ctypes.byref(md)))
        bottom_degrees = int(bottom_percentile * md.value)              # networkx_graph.py:60
        nn, ne = ctypes.c_uint64(), ctypes.c_uint64()
        check(self.lib.pdg_filter_select(self.h, bottom_degrees, ctypes.byref(nn), ctypes.byref(ne)))
        n = ne.value
        src, tgt = np.zeros(n, dtype=np.uint64), np.zeros(n, dtype=np.uint64)
        ty, val = np.zeros(n, dtype=np.uint32), np.zeros(n, dtype=np.int32)
        check(self.lib.pdg_filter_fetch(self.h, src.ctypes.data, tgt.ctypes.data, ty.ctypes.data, val.ctypes.data))
        return nn.value, src, tgt, ty, val

    # ---- multi GPU ----
    def comm_unique_id(self):
        buf = np.zeros(128, dtype=np.uint8)
        check(self.lib.pdg_comm_unique_id(buf.ctypes.data))
        return buf

    def comm_init(self, uid, rank, world):
        uid = np.ascontiguousarray(uid, dtype=np.uint8)
        check(self.lib.pdg_comm_init(self.h, uid.ctypes.data, int(rank), int(world)))
